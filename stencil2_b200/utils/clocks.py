"""Sample SM clocks and throttle reasons WHILE a benchmark's timed region runs (B200_PROFILING.md, clocks line).
Uses NVML through pynvml when available, else `nvidia-smi --query-gpu ... -lms`. Never changes clocks."""
import shutil
import statistics
import subprocess
import threading
import time

_REASONS = {
    0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x10: "sync_boost",
    0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown", 0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting",
}


class ClockSampler:
    def __init__(self, device_index=0, period_s=0.05, busy_only=False):
        """busy_only: keep only samples taken while the GPU was not idle (NVML's gpu_idle reason clear) — for a timed
        region that this process cannot bracket exactly, e.g. a benchmark binary run as a subprocess, whose set-up
        phase would otherwise drag the median down to the idle clock."""
        self.idx, self.period, self.busy_only = device_index, period_s, busy_only
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self.idle_samples = []
        self._stop = threading.Event()
        self._thread = None
        self._proc = None
        self._mode = None

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            self._nv = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(self.idx)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM)
            self._mode = "nvml"
            self._thread = threading.Thread(target=self._loop, daemon=True)
            self._thread.start()
        except Exception:
            smi = shutil.which("nvidia-smi")
            if smi:
                self._mode = "smi"
                q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
                    "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
                self._proc = subprocess.Popen([smi, f"--id={self.idx}", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                              stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        return self

    def _loop(self):
        nv = self._nv
        while not self._stop.is_set():
            try:
                mhz = nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_SM)
                try:
                    mask = nv.nvmlDeviceGetCurrentClocksEventReasons(self._h)
                except Exception:
                    mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self._h)
                if self.busy_only and (mask & 0x1):
                    self.idle_samples.append(mhz)
                    time.sleep(self.period)
                    continue
                self.samples.append(mhz)
                for bit, name in _REASONS.items():
                    if mask & bit and name != "gpu_idle":
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(self.period)

    def stop(self):
        self._stop.set()
        if self._thread:
            self._thread.join(timeout=2)
        if self._proc:
            self._proc.terminate()
            try:
                out, _ = self._proc.communicate(timeout=2)
            except Exception:
                out = ""
            names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
            for line in out.splitlines():
                f = [t.strip() for t in line.split(",")]
                try:
                    self.samples.append(int(float(f[0])))
                    self.max_mhz = int(float(f[1]))
                    for n, v in zip(names, f[2:6]):
                        if v.lower().startswith("active"):
                            self.reasons.add(n)
                except Exception:
                    continue
        return self.summary()

    def summary(self):
        if not self.samples and self.idle_samples:  # never saw the GPU busy: say so instead of reporting nothing
            return {"sm_mhz": int(statistics.median(self.idle_samples)), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                    "samples": 0, "idle_samples": len(self.idle_samples), "source": self._mode, "note": "no sample caught the GPU busy"}
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "source": self._mode or "unavailable"}
        return {"sm_mhz": int(statistics.median(self.samples)), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples), "source": self._mode}
