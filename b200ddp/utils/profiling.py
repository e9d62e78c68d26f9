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
    """Samples SM clocks / throttle reasons of one GPU: NVML polled from a thread when ``pynvml``
    is importable (sub-millisecond period, and ``sample_now()`` for timed regions shorter than
    any period), else ``nvidia-smi -lms`` in a child process."""

    _REASONS = (("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20),
                ("sw_power_cap", 0x4))

    def __init__(self, index: int = 0, period_ms: int = 100):
        self.index = index
        self.period_ms = period_ms
        self.rows: List[List[str]] = []
        self._proc = None
        self._thread = None
        self._nvml = None
        self._handle = None
        self._stop = threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self._handle = pynvml.nvmlDeviceGetHandleByIndex(self._physical_index(index))
            self._nvml = pynvml
        except Exception:  # noqa: BLE001 - no NVML on this box: nvidia-smi fallback
            self._nvml = None

    @staticmethod
    def _physical_index(index: int) -> int:
        import os
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        if vis:
            ids = [v.strip() for v in vis.split(",") if v.strip()]
            if index < len(ids) and ids[index].isdigit():
                return int(ids[index])
        return index

    def sample_now(self) -> None:
        """One NVML reading, appended to the record (call while the GPU is busy)."""
        nv = self._nvml
        if nv is None:
            return
        try:
            sm = nv.nvmlDeviceGetClockInfo(self._handle, nv.NVML_CLOCK_SM)
            mx = nv.nvmlDeviceGetMaxClockInfo(self._handle, nv.NVML_CLOCK_SM)
            try:
                mask = nv.nvmlDeviceGetCurrentClocksEventReasons(self._handle)
            except Exception:  # noqa: BLE001 - older binding name
                mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self._handle)
            try:
                power = nv.nvmlDeviceGetPowerUsage(self._handle) / 1000.0
            except Exception:  # noqa: BLE001
                power = 0.0
            self.rows.append([str(sm), str(mx), f"{power:.1f}"] +
                             ["Active" if mask & bit else "Not Active" for _, bit in self._REASONS])
        except Exception:  # noqa: BLE001 - a failed reading is a missing sample, not an error
            pass

    def start(self):
        if self._nvml is not None:
            def poll():
                while not self._stop.is_set():
                    self.sample_now()
                    self._stop.wait(self.period_ms / 1000.0)
            self._thread = threading.Thread(target=poll, daemon=True)
            self._thread.start()
            return self
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
        self._stop.set()
        if self._nvml is not None and self._thread is not None:
            self._thread.join(timeout=2)
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
