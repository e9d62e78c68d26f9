"""Step timing, NVTX ranges and clock sampling.

The reference only has wall-clock ``time.time()`` around an epoch (reference
``train.py:262-312``).  Benchmarks here are device-timed with CUDA events, taken
as the max over ranks, with SM clocks and throttle reasons sampled during the
timed region (see ``/bench.py``).
"""
from __future__ import annotations

import contextlib
import statistics
import subprocess
import threading
from typing import Dict, List, Optional

import torch

_NVTX = False


def enable_nvtx(on: bool = True) -> None:
    global _NVTX
    _NVTX = on and torch.cuda.is_available()


@contextlib.contextmanager
def nvtx_range(name: str):
    if _NVTX:
        torch.cuda.nvtx.range_push(name)
        try:
            yield
        finally:
            torch.cuda.nvtx.range_pop()
    else:
        yield


class DeviceTimer:
    """CUDA-event bracket on the current stream; ``ms`` is valid after ``stop``."""

    def __init__(self):
        self.t0 = torch.cuda.Event(enable_timing=True)
        self.t1 = torch.cuda.Event(enable_timing=True)
        self.ms: Optional[float] = None

    def start(self):
        torch.cuda.synchronize()
        self.t0.record()

    def stop(self) -> float:
        self.t1.record()
        torch.cuda.synchronize()
        self.ms = self.t0.elapsed_time(self.t1)
        return self.ms


_QUERY = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
          "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
          "clocks_event_reasons.sw_power_cap")


class ClockSampler:
    """Samples ``nvidia-smi`` clocks/throttle reasons for one GPU in a thread."""

    def __init__(self, index: int = 0, period_ms: int = 100):
        self.index = index
        self.period_ms = period_ms
        self.rows: List[List[str]] = []
        self._proc = None
        self._thread = None

    def start(self):
        try:
            self._proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.index}", f"--query-gpu={_QUERY}",
                 "--format=csv,noheader,nounits", "-lms", str(self.period_ms)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self._proc = None
            return self

        def pump():
            for line in self._proc.stdout:
                parts = [p.strip() for p in line.split(",")]
                if len(parts) >= 7:
                    self.rows.append(parts)

        self._thread = threading.Thread(target=pump, daemon=True)
        self._thread.start()
        return self

    def stop(self) -> Dict:
        if self._proc is not None:
            self._proc.terminate()
            try:
                self._proc.wait(timeout=2)
            except subprocess.TimeoutExpired:
                self._proc.kill()
            if self._thread is not None:
                self._thread.join(timeout=2)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
            except ValueError:
                continue
            for name, val in zip(names, r[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {
            "sm_mhz": statistics.median(sm) if sm else None,
            "sm_max_mhz": max(mx) if mx else None,
            "reasons": sorted(reasons),
            "samples": len(sm),
        }
