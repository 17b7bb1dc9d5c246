"""Device-side timing and clock sampling.

The reference's only instrumentation is a host wall-clock accumulator over iterations 1-39
(/root/reference/part1/main.py:66,86-91).  On a GPU that would time the
launch queue, so benchmarks here use CUDA events on the launching stream, take the maximum
over ranks, and sample SM clocks / throttle reasons with nvidia-smi while the timed region
runs (B200_PROFILING.md "clocks DURING the timed region").
"""
from __future__ import annotations

import statistics
import subprocess
import threading
import time

import torch

_QUERY = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
          "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
          "clocks_event_reasons.sw_power_cap")


class ClockSampler:
    """Background nvidia-smi poller; ``summary()`` gives the bench.py "clocks" object."""

    def __init__(self, gpu_index: int = 0, period_s: float = 0.2):
        self.gpu_index, self.period_s = gpu_index, period_s
        self.samples: list[list[str]] = []
        self._stop = threading.Event()
        self._thread = None

    def _poll(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--id={self.gpu_index}", f"--query-gpu={_QUERY}",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5)
                if out.returncode == 0 and out.stdout.strip():
                    self.samples.append([f.strip() for f in out.stdout.strip().splitlines()[0].split(",")])
            except Exception:  # noqa: BLE001 - sampling must never break a benchmark
                pass
            self._stop.wait(self.period_s)

    def __enter__(self):
        self._thread = threading.Thread(target=self._poll, daemon=True)
        self._thread.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        if self._thread is not None:
            self._thread.join(timeout=6)

    def summary(self) -> dict:
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for s in self.samples:
            try:
                sm.append(float(s[0]))
                mx.append(float(s[1]))
                pw.append(float(s[2]))
            except (ValueError, IndexError):
                continue
            for name, val in zip(names, s[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


class L2Flusher:
    """Writes a buffer larger than the 126 MB L2 so the next timed iteration starts cold."""

    def __init__(self, device, mbytes: int = 256):
        self.buf = torch.empty(mbytes << 20, dtype=torch.uint8, device=device)

    def flush(self):
        self.buf.zero_()


def max_over_ranks(value: float, device) -> float:
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return value
    t = torch.tensor([value], dtype=torch.float64, device=device if dist.get_backend() == "nccl" else "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def wall_ns() -> int:
    return time.perf_counter_ns()
