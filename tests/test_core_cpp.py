"""Runs the C++ host-runtime test binary (csrc/tests/core_tests.cpp): the reference's 50 gtest
cases restated against this implementation plus the cases the reference lacks.  CPU only."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "build", "bin", "core_tests")


def _build():
    sys.path.insert(0, ROOT)
    import importlib.util
    spec = importlib.util.spec_from_file_location("minips_build", os.path.join(ROOT, "build.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    mod.build_core(verbose=False)


@pytest.fixture(scope="module")
def binary():
    _build()
    assert os.path.exists(BIN)
    return BIN


def test_all_cpp_cases(binary):
    p = subprocess.run([binary], capture_output=True, text=True, timeout=600, errors="replace")
    tail = "\n".join(p.stdout.splitlines()[-15:])
    failed = [l for l in p.stdout.splitlines() if l.startswith("[ FAIL ]")]
    assert p.returncode == 0 and not failed, f"{failed}\n{tail}\n{p.stderr[-3000:]}"
    last = p.stdout.strip().splitlines()[-1]
    n = int(last.split()[0])
    assert n >= 70, last  # 50 reference cases + the extra coverage


def test_reference_case_names_present(binary):
    names = subprocess.run([binary, "--list"], capture_output=True, text=True).stdout.split()
    for must in ["RangePartitionManager.SliceKeys", "SimpleIdMapper.AllocateDeallocateThread",
                 "WorkerSpec.InsertWorkerIdThreadId", "MapStorage.SubAddSubGet", "PendingBuffer.PushAndPop",
                 "ProgressTracker.Advance", "ServerThread.Get", "BSPModel.CheckGetAndAdd",
                 "SSPModel.CheckStaleness", "ASPModel.CheckGetAndAdd", "Sender.Send", "Mailbox.BarrierFourNodes",
                 "KVClientTable.Get", "Engine.MultipleTasks", "Engine.KVClientTableMapStorage"]:
        assert must in names
