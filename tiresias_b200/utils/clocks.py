"""SM-clock / throttle-reason sampling during a timed region (B200 profiling recipe: a number is
only valid with the clocks seen while it was measured)."""
from __future__ import annotations

import statistics
import threading
import time
from typing import Dict, List, Optional

_REASONS = {
    0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
    0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
    0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting",
}
REJECT = {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}


class ClockSampler:
    """Polls NVML every ``period`` seconds on a background thread between start() and stop()."""

    def __init__(self, device_index: int = 0, period: float = 0.1):
        self.index, self.period = device_index, period
        self._samples: List[int] = []
        self._reasons: set = set()
        self._max = 0
        self._stop = threading.Event()
        self._thread: Optional[threading.Thread] = None
        self._nvml = None
        self._handle = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self._nvml = pynvml
            self._handle = pynvml.nvmlDeviceGetHandleByIndex(device_index)
            self._max = int(pynvml.nvmlDeviceGetMaxClockInfo(self._handle, pynvml.NVML_CLOCK_SM))
        except Exception:
            self._nvml = None

    def _loop(self) -> None:
        nv = self._nvml
        while not self._stop.is_set():
            try:
                self._samples.append(int(nv.nvmlDeviceGetClockInfo(self._handle, nv.NVML_CLOCK_SM)))
                mask = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self._handle))
                for bit, name in _REASONS.items():
                    if mask & bit and name != "gpu_idle":
                        self._reasons.add(name)
            except Exception:
                pass
            time.sleep(self.period)

    def start(self) -> "ClockSampler":
        if self._nvml is not None:
            self._thread = threading.Thread(target=self._loop, daemon=True)
            self._thread.start()
        return self

    def stop(self) -> Dict[str, object]:
        self._stop.set()
        if self._thread is not None:
            self._thread.join(timeout=2.0)
        med = int(statistics.median(self._samples)) if self._samples else None
        return {"sm_mhz": med, "sm_max_mhz": self._max or None, "reasons": sorted(self._reasons),
                "samples": len(self._samples)}


def rejected(clocks: Dict[str, object]) -> bool:
    return bool(REJECT.intersection(clocks.get("reasons") or []))


class UtilSampler:
    """GPU utilisation of one device while a job runs (K7): NVML's utilisation rate — the share of the
    last sampling window in which at least one kernel was resident, the quantity the trace columns
    ``gpu_utilization_avg`` / ``gpu_utilization_max`` report — polled on a background thread. The
    reference draws this number from N(avg, (max-avg)/2) per task and tick (infra/device.py:30-36);
    calibrated mode replaces the draw with what stop() returns. No NVML (CPU loopback) -> ``None``s."""

    def __init__(self, device_index: int = 0, period: float = 0.05):
        self.index, self.period = device_index, period
        self._samples: List[int] = []
        self._stop = threading.Event()
        self._thread: Optional[threading.Thread] = None
        self._nvml = None
        self._handle = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self._nvml = pynvml
            self._handle = pynvml.nvmlDeviceGetHandleByIndex(device_index)
        except Exception:
            self._nvml = None

    def _loop(self) -> None:
        nv = self._nvml
        while not self._stop.is_set():
            try:
                self._samples.append(int(nv.nvmlDeviceGetUtilizationRates(self._handle).gpu))
            except Exception:
                pass
            time.sleep(self.period)

    def start(self) -> "UtilSampler":
        if self._nvml is not None:
            self._thread = threading.Thread(target=self._loop, daemon=True)
            self._thread.start()
        return self

    def stop(self) -> Dict[str, Optional[float]]:
        self._stop.set()
        if self._thread is not None:
            self._thread.join(timeout=2.0)
        if not self._samples:
            return {"gpu_util_avg": None, "gpu_util_max": None, "samples": 0}
        return {"gpu_util_avg": sum(self._samples) / len(self._samples), "gpu_util_max": float(max(self._samples)),
                "samples": len(self._samples)}
