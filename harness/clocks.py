"""SM clock / throttle-reason sampling for the timed regions of bench.py."""
import os
import threading
import time


class ClockSampler:
    """SM clock / throttle-reason sampling while a timed region runs (B200_PROFILING.md): NVML polled from
    a thread every 5 ms (nvidia-smi -lms cannot sample a region shorter than its period)."""

    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}

    def __init__(self, index):
        self.index = index
        self.samples = []
        self.stop_flag = threading.Event()
        self.thread = None
        self.handle = None
        self.sm_max = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            # LOCAL_RANK indexes the visible devices; NVML enumerates all of them
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[index]) if vis and vis.split(",")[index].isdigit() else index
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.sm_max = float(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.handle = None

    def _poll(self):
        nv = self.nv
        reasons = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or \
            nv.nvmlDeviceGetCurrentClocksThrottleReasons
        while not self.stop_flag.is_set():
            try:
                self.samples.append((float(nv.nvmlDeviceGetClockInfo(self.handle, nv.NVML_CLOCK_SM)),
                                     int(reasons(self.handle))))
            except Exception:
                pass
            time.sleep(0.005)

    def start(self):
        if self.handle is None:
            return
        self.thread = threading.Thread(target=self._poll, daemon=True)
        self.thread.start()

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": self.sm_max, "reasons": [], "samples": 0}
        if self.thread is None:
            return out
        self.stop_flag.set()
        self.thread.join(timeout=2)
        if self.samples:
            sm = sorted(x[0] for x in self.samples)
            mask = 0
            for _, m in self.samples:
                mask |= m
            out.update(sm_mhz=sm[len(sm) // 2], reasons=sorted(n for b, n in self.REASONS.items() if mask & b),
                       samples=len(sm))
        return out
