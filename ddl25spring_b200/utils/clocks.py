"""Sample SM clocks and throttle reasons DURING a timed region (B200_PROFILING.md recipe)."""
from __future__ import annotations

import statistics
import subprocess
import threading
import time

_REASON_BITS = {
    0x0000000000000004: "sw_power_cap",
    0x0000000000000008: "hw_slowdown",
    0x0000000000000020: "sw_thermal_slowdown",
    0x0000000000000040: "hw_thermal_slowdown",
    0x0000000000000080: "hw_power_brake_slowdown",
}


class ClockSampler:
    """Background sampler: NVML when importable, else ``nvidia-smi`` polling."""

    def __init__(self, device_index: int = 0, period_s: float = 0.02):
        self.idx, self.period = device_index, period_s
        self.samples: list[int] = []
        self.reasons: set[str] = set()
        self.sm_max = None
        self._stop = threading.Event()
        self._thr: threading.Thread | None = None
        self._nvml = None
        try:
            import pynvml

            pynvml.nvmlInit()
            self._nvml = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(device_index)
            self.sm_max = int(pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self._nvml = None

    def _poll_nvml(self):
        n = self._nvml
        self.samples.append(int(n.nvmlDeviceGetClockInfo(self._h, n.NVML_CLOCK_SM)))
        try:
            mask = int(n.nvmlDeviceGetCurrentClocksEventReasons(self._h))
        except Exception:
            mask = int(n.nvmlDeviceGetCurrentClocksThrottleReasons(self._h))
        for bit, name in _REASON_BITS.items():
            if mask & bit:
                self.reasons.add(name)

    def _poll_smi(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        out = subprocess.run(["nvidia-smi", f"--id={self.idx}", f"--query-gpu={q}", "--format=csv,noheader,nounits"],
                             capture_output=True, text=True, timeout=5).stdout.strip().split(",")
        self.samples.append(int(float(out[0])))
        self.sm_max = int(float(out[1]))
        for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), out[2:]):
            if "Active" in v and "Not" not in v:
                self.reasons.add(name)

    def _loop(self):
        while not self._stop.is_set():
            try:
                (self._poll_nvml if self._nvml else self._poll_smi)()
            except Exception:
                pass
            time.sleep(self.period)

    def __enter__(self):
        self._thr = threading.Thread(target=self._loop, daemon=True)
        self._thr.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        if self._thr:
            self._thr.join(timeout=5)

    def summary(self) -> dict:
        return {"sm_mhz": (statistics.median(self.samples) if self.samples else None), "sm_max_mhz": self.sm_max,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}
