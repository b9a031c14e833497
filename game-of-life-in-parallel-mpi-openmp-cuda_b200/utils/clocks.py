"""SM clock / throttle sampling during a timed region (B200 profiling recipe, clocks line)."""
from __future__ import annotations

import shutil
import statistics
import subprocess
import threading
import time

_QUERY = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
          "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
          "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")


class ClockSampler:
    """Samples `nvidia-smi` every `period_s` on a background thread between start() and stop()."""

    def __init__(self, gpu_index: int = 0, period_s: float = 0.1):
        self.gpu, self.period = gpu_index, period_s
        self.samples: list[list[str]] = []
        self._stop = threading.Event()
        self._thread = None

    def _poll_nvml(self) -> bool:
        """Fast path: NVML through nvidia-ml-py (one sample per ~2 ms instead of ~80 ms per nvidia-smi call)."""
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.gpu)
            mx = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
        except Exception:
            return False
        bits = {"hw_slowdown": nv.nvmlClocksThrottleReasonHwSlowdown,
                "hw_thermal_slowdown": nv.nvmlClocksThrottleReasonHwThermalSlowdown,
                "sw_thermal_slowdown": nv.nvmlClocksThrottleReasonSwThermalSlowdown,
                "sw_power_cap": nv.nvmlClocksThrottleReasonSwPowerCap}
        while not self._stop.is_set():
            try:
                sm = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
                pw = nv.nvmlDeviceGetPowerUsage(h) / 1000.0
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                flags = ["Active" if (r & b) else "Not Active" for b in bits.values()]
                self.samples.append([str(sm), str(mx), f"{pw:.1f}", hex(r)] + flags)
            except Exception:
                pass
            self._stop.wait(min(self.period, 0.005))
        return True

    def _poll(self):
        if self._poll_nvml():
            return
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--id={self.gpu}", f"--query-gpu={_QUERY}",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5)
                if out.returncode == 0 and out.stdout.strip():
                    self.samples.append([f.strip() for f in out.stdout.strip().splitlines()[0].split(",")])
            except Exception:
                pass
            self._stop.wait(self.period)

    def start(self):
        self._thread = threading.Thread(target=self._poll, daemon=True)
        self._thread.start()
        return self

    def stop(self) -> dict:
        self._stop.set()
        if self._thread is not None:
            self._thread.join(timeout=10)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        sm = [float(s[0]) for s in self.samples if s[0].replace(".", "").isdigit()]
        mx = [float(s[1]) for s in self.samples if s[1].replace(".", "").isdigit()]
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for s in self.samples:
            for name, val in zip(names, s[4:8]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        pw = []
        for s in self.samples:
            try:
                pw.append(float(s[2]))
            except ValueError:
                pass
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(self.samples), "power_w_max": max(pw) if pw else None}


class Timer:
    """Wall-clock phase timer (reading / loop / writing lines of the MPI programs, M:187,262..)."""

    def __init__(self):
        self.t0 = time.perf_counter()

    def ms(self) -> float:
        return (time.perf_counter() - self.t0) * 1e3
