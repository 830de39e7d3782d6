"""`bench.py --impl reference`: the unmodified reference built under the shim headers runs the bench
config (scaled down) through its own LRExample and reports the contract's JSON line."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_BIN = os.path.join(ROOT, "baseline", "_ref", "bin", "LRExample")


@pytest.mark.skipif(not os.path.exists(REF_BIN),
    reason="reference shim build not present (baseline/ref_build/build_ref.py)")
def test_reference_arm_small_run(tmp_path):
    env = dict(os.environ, MINIPS_REF_DATA_DIR=str(tmp_path), MINIPS_REF_WORKERS_PER_NODE="1",
               MINIPS_REF_TIMEOUT="150")
    env.pop("RANK", None)
    env.pop("WORLD_SIZE", None)
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "5",
                        "--warmup", "3", "--samples", "200000", "--features", "100000", "--batch", "4096"],
                       capture_output=True, text=True, timeout=240, env=env, cwd=ROOT)
    assert p.returncode == 0, p.stderr[-2000:]
    rec = json.loads(p.stdout.strip().splitlines()[-1])
    assert rec["impl"] == "reference"
    assert "unavailable" not in rec, rec
    assert rec["metric"] == "sparse_lr_samples_per_sec" and rec["value"] > 0
    assert rec["config"]["server_shards"] == 8 and rec["config"]["global_batch"] == 4096
    assert rec["steps"] == 5 and rec["gpu_launches"] == 0


def test_reference_arm_reports_unavailable_without_build(tmp_path, monkeypatch):
    sys.path.insert(0, os.path.join(ROOT, "baseline"))
    import ref_run
    monkeypatch.setattr(ref_run, "REF_BIN", str(tmp_path / "missing"))
    monkeypatch.delenv("RANK", raising=False)

    class A:
        pass
    rec = ref_run.run(A())
    assert rec["impl"] == "reference" and "unavailable" in rec
