"""Sample SM clocks and throttle reasons while a timed region runs (B200_PROFILING.md
'clocks DURING the timed region'). Uses NVML in a background thread; falls back to
one `nvidia-smi` query when NVML is unavailable."""
from __future__ import annotations

import statistics
import subprocess
import threading
import time

_REASONS = {
    0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
    0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown", 0x80: "hw_power_brake_slowdown",
    0x100: "display_clock_setting",
}


class ClockSampler:
    def __init__(self, device: int = 0, period_s: float = 0.05):
        self.device, self.period = device, period_s
        self.samples, self.reasons = [], set()
        self.sm_max = None
        self._stop = threading.Event()
        self._thr = None
        self._h = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self._nv = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(device)
            self.sm_max = pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self._nv = None

    def _once(self):
        if self._nv is None:
            return
        try:
            self.samples.append(self._nv.nvmlDeviceGetClockInfo(self._h, self._nv.NVML_CLOCK_SM))
            mask = self._nv.nvmlDeviceGetCurrentClocksEventReasons(self._h) if hasattr(
                self._nv, "nvmlDeviceGetCurrentClocksEventReasons") else self._nv.nvmlDeviceGetCurrentClocksThrottleReasons(self._h)
            for bit, name in _REASONS.items():
                if mask & bit and name != "gpu_idle":
                    self.reasons.add(name)
        except Exception:
            pass

    def _run(self):
        while not self._stop.is_set():
            self._once()
            time.sleep(self.period)

    def __enter__(self):
        if self._nv is not None:
            self._thr = threading.Thread(target=self._run, daemon=True)
            self._thr.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        if self._thr:
            self._thr.join(timeout=2)
        if self._nv is not None:
            self._once()

    def summary(self) -> dict:
        if not self.samples:
            try:  # single nvidia-smi sample as a last resort
                out = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,clocks.max.sm", "--format=csv,noheader,nounits",
                                      "-i", str(self.device)], capture_output=True, text=True, timeout=10).stdout
                a, b = [int(x) for x in out.strip().split(",")]
                return {"sm_mhz": a, "sm_max_mhz": b, "reasons": [], "samples": 1, "source": "nvidia-smi"}
            except Exception:
                return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0, "source": "none"}
        return {"sm_mhz": int(statistics.median(self.samples)), "sm_max_mhz": self.sm_max,
                "reasons": sorted(self.reasons), "samples": len(self.samples), "source": "nvml"}
