"""Benchmark hygiene helpers: device timing (max over ranks), nvidia-smi clock sampling
during the timed region, L2 flush buffer."""
from __future__ import annotations

import os
import shutil
import statistics
import subprocess
import tempfile
import time
from typing import Dict, List, Optional

import torch

_Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
      "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
      "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")


class ClockSampler:
    """Samples SM clocks / throttle reasons with nvidia-smi while a timed region runs."""

    def __init__(self, device_index: int = 0, period_ms: int = 100):
        self.device_index = device_index
        self.period_ms = period_ms
        self.proc: Optional[subprocess.Popen] = None
        self.path: Optional[str] = None

    def start(self):
        exe = shutil.which("nvidia-smi")
        if exe is None:
            return self
        fd, self.path = tempfile.mkstemp(prefix="mps_clocks_", suffix=".csv")
        os.close(fd)
        self._out = open(self.path, "w")
        phys = os.environ.get("CUDA_VISIBLE_DEVICES")
        idx = self.device_index
        if phys:
            try:
                idx = int(phys.split(",")[self.device_index])
            except Exception:
                pass
        self.proc = subprocess.Popen(
            [exe, f"--id={idx}", f"--query-gpu={_Q}", "--format=csv,noheader,nounits", "-lms", str(self.period_ms)],
            stdout=self._out, stderr=subprocess.DEVNULL)
        # nvidia-smi needs a few hundred ms to come up (NVML init); wait for its first line so that its start-up
        # does not land inside the first timed region
        end = time.time() + 2.0
        while time.time() < end and os.path.getsize(self.path) == 0 and self.proc.poll() is None:
            time.sleep(0.02)
        return self

    def stop(self) -> Dict:
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        time.sleep(self.period_ms / 1000.0)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        self._out.close()
        sm: List[float] = []
        mx: List[float] = []
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        try:
            for line in open(self.path):
                f = [x.strip() for x in line.split(",")]
                if len(f) < 9:
                    continue
                try:
                    sm.append(float(f[1]))
                    mx.append(float(f[2]))
                except ValueError:
                    continue
                for n, v in zip(names, f[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
        finally:
            try:
                os.unlink(self.path)
            except OSError:
                pass
        if sm:
            # "under load": the upper half of the samples (idle gaps between launches excluded)
            hi = sorted(sm)[len(sm) // 2:]
            out["sm_mhz"] = statistics.median(hi)
            out["sm_max_mhz"] = max(mx)
            out["samples"] = len(sm)
        out["reasons"] = sorted(reasons)
        return out


def max_over_ranks(value: float, device) -> float:
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return value
    t = torch.tensor([value], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sync_barrier(device):
    import torch.distributed as dist
    torch.cuda.synchronize(device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.barrier()
    torch.cuda.synchronize(device)


class L2Flusher:
    """Writes a buffer larger than the 126 MB L2 between timed iterations."""

    def __init__(self, device, nbytes: int = 256 << 20):
        self.buf = torch.empty(nbytes // 4, dtype=torch.int32, device=device)

    def flush(self):
        self.buf.add_(1)
