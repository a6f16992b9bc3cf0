"""Sample SM clocks and throttle reasons while a timed region runs (bench.py contract)."""
from __future__ import annotations

import statistics
import subprocess
import threading
import time
from typing import Dict, List, Optional

_REASONS = {
    0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
    0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
    0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting",
}


class ClockSampler:
    """Background thread polling NVML (or nvidia-smi when pynvml is unavailable)."""

    def __init__(self, index: int = 0, period_s: float = 0.02):
        self.index, self.period = index, period_s
        self.sm: List[int] = []
        self.reasons = set()
        self.max_mhz: Optional[int] = None
        self._stop = threading.Event()
        self._thr: Optional[threading.Thread] = None
        self._nvml = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self._nvml = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self._nvml = None

    def _sample_once(self) -> None:
        if self._nvml is not None:
            n = self._nvml
            self.sm.append(int(n.nvmlDeviceGetClockInfo(self._h, n.NVML_CLOCK_SM)))
            try:
                mask = int(n.nvmlDeviceGetCurrentClocksEventReasons(self._h))
            except Exception:
                mask = int(n.nvmlDeviceGetCurrentClocksThrottleReasons(self._h))
            for bit, name in _REASONS.items():
                if mask & bit and name != "gpu_idle":
                    self.reasons.add(name)
        else:
            q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
            out = subprocess.run(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}", "--format=csv,noheader,nounits"],
                                 capture_output=True, text=True, timeout=5).stdout.strip().split(",")
            self.sm.append(int(float(out[0])))
            self.max_mhz = int(float(out[1]))
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), out[2:]):
                if "Active" in val and "Not" not in val:
                    self.reasons.add(name)

    def _loop(self) -> None:
        while not self._stop.is_set():
            try:
                self._sample_once()
            except Exception:
                pass
            time.sleep(self.period)

    def __enter__(self) -> "ClockSampler":
        self._thr = threading.Thread(target=self._loop, daemon=True)
        self._thr.start()
        return self

    def __exit__(self, *exc) -> None:
        self._stop.set()
        if self._thr is not None:
            self._thr.join(timeout=5)

    def summary(self) -> Dict:
        return {"sm_mhz": (statistics.median(self.sm) if self.sm else None), "sm_max_mhz": self.max_mhz,
                "samples": len(self.sm), "reasons": sorted(self.reasons)}
