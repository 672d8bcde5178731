"""SM-clock / throttle-reason sampling during a timed region (B200_PROFILING.md "clocks line").

Uses NVML (nvidia_ml_py) in a background thread; falls back to one ``nvidia-smi`` query.  The
sampler never changes clocks — it only reads them."""
from __future__ import annotations

import statistics
import subprocess
import threading
import time
from typing import List, Optional

_REASONS = [  # (name, NVML bit)
    ("gpu_idle", 0x1), ("applications_clocks_setting", 0x2), ("sw_power_cap", 0x4),
    ("hw_slowdown", 0x8), ("sync_boost", 0x10), ("sw_thermal_slowdown", 0x20),
    ("hw_thermal_slowdown", 0x40), ("hw_power_brake_slowdown", 0x80), ("display_clock_setting", 0x100),
]


class ClockSampler:
    def __init__(self, device_index: int = 0, period_s: float = 0.05):
        self.idx, self.period = device_index, period_s
        self.samples: List[int] = []
        self.reasons = set()
        self.sm_max: Optional[int] = None
        self._stop = threading.Event()
        self._thr: Optional[threading.Thread] = None
        self._nvml = None

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            self._nvml = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(self.idx)
            self.sm_max = int(pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self._nvml = None
        self._thr = threading.Thread(target=self._run, daemon=True)
        self._thr.start()
        return self

    def _sample_nvml(self):
        n = self._nvml
        self.samples.append(int(n.nvmlDeviceGetClockInfo(self._h, n.NVML_CLOCK_SM)))
        try:
            bits = int(n.nvmlDeviceGetCurrentClocksEventReasons(self._h))
        except Exception:
            bits = int(n.nvmlDeviceGetCurrentClocksThrottleReasons(self._h))
        for name, b in _REASONS:
            if bits & b and name != "gpu_idle":
                self.reasons.add(name)

    def _sample_smi(self):
        q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown," \
            "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown," \
            "clocks_event_reasons.sw_power_cap"
        out = subprocess.run(["nvidia-smi", f"--id={self.idx}", f"--query-gpu={q}",
                              "--format=csv,noheader,nounits"], capture_output=True, text=True,
                             timeout=10).stdout.strip().split(",")
        self.samples.append(int(out[0]))
        self.sm_max = int(out[1])
        for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown",
                            "sw_power_cap"), out[2:]):
            if v.strip().lower().startswith("active"):
                self.reasons.add(name)

    def _run(self):
        while not self._stop.is_set():
            try:
                if self._nvml is not None:
                    self._sample_nvml()
                else:
                    self._sample_smi()
            except Exception:
                pass
            self._stop.wait(self.period)

    def stop(self) -> dict:
        self._stop.set()
        if self._thr:
            self._thr.join(timeout=5)
        med = int(statistics.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.sm_max, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


def rejected(clocks: dict) -> bool:
    """A timed region must be measured again when the GPU was slowed down by hardware or thermals, or when its SM clock
    sat well below the maximum for no stated reason (a leftover clock lock).  A software power cap is kept and
    reported: that is the card's steady state under this load."""
    reasons = set(clocks.get("reasons", []))
    if reasons & {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}:
        return True
    sm, top = clocks.get("sm_mhz"), clocks.get("sm_max_mhz")
    if sm and top and sm < 0.8 * top and not (reasons & {"sw_power_cap", "hw_power_brake_slowdown"}):
        return True
    return False
