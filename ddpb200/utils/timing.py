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
    """Background poller of SM clock, power and throttle reasons; ``summary()`` gives the bench.py
    "clocks" object.  Uses NVML in-process when available (about a millisecond per sample, so a
    short timed region still gets tens of samples); falls back to spawning nvidia-smi."""

    # nvmlClocksThrottleReason* / nvmlClocksEventReason* bits
    _REASON_BITS = {"sw_power_cap": 0x4, "hw_slowdown": 0x8, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40}

    def __init__(self, gpu_index: int = 0, period_s: float = 0.01, uuid: str | None = None):
        self.gpu_index, self.period_s, self.uuid = gpu_index, period_s, uuid
        self.samples: list[list[str]] = []
        self._stop = threading.Event()
        self._thread = None
        self.source = "nvidia-smi"

    def _nvml_handle(self):
        try:
            import pynvml

            pynvml.nvmlInit()
            if self.uuid:
                try:
                    u = self.uuid if str(self.uuid).startswith("GPU-") else f"GPU-{self.uuid}"
                    return pynvml, pynvml.nvmlDeviceGetHandleByUUID(u.encode() if hasattr(u, "encode") else u)
                except Exception:  # noqa: BLE001
                    pass
            return pynvml, pynvml.nvmlDeviceGetHandleByIndex(self.gpu_index)
        except Exception:  # noqa: BLE001 - no NVML: use the command line tool
            return None, None

    def _sample_nvml(self, nv, h) -> bool:
        try:
            sm = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
            mx = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            pw = nv.nvmlDeviceGetPowerUsage(h) / 1000.0
            get = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or \
                getattr(nv, "nvmlDeviceGetCurrentClocksThrottleReasons")
            mask = int(get(h))
            names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
            flags = ["Active" if mask & self._REASON_BITS[n] else "Not Active" for n in names]
            self.samples.append([str(sm), str(mx), str(pw), *flags])
            return True
        except Exception:  # noqa: BLE001
            return False

    def _sample_smi(self):
        try:
            out = subprocess.run(["nvidia-smi", f"--id={self.gpu_index}", f"--query-gpu={_QUERY}",
                                  "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5)
            if out.returncode == 0 and out.stdout.strip():
                self.samples.append([f.strip() for f in out.stdout.strip().splitlines()[0].split(",")])
        except Exception:  # noqa: BLE001 - sampling must never break a benchmark
            pass

    def _poll(self):
        nv, h = self._nvml_handle()
        if nv is not None:
            self.source = "nvml"
        while not self._stop.is_set():
            if nv is None or not self._sample_nvml(nv, h):
                nv = None
                self.source = "nvidia-smi"
                self._sample_smi()
                self._stop.wait(max(self.period_s, 0.1))
            else:
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
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "source": self.source,
                "reasons": sorted(reasons)}


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
