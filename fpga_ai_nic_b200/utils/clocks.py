"""Sample SM clocks and throttle reasons while a timed region runs (B200_PROFILING.md clocks line).

Primary path: NVML in a background thread (a few ms per sample, so even a 20 ms timed region gets
several samples).  Fallback: an ``nvidia-smi -lms`` child process.
"""
from __future__ import annotations

import os
import shutil
import statistics
import subprocess
import tempfile
import threading
import time

_QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
          "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
          "clocks_event_reasons.sw_power_cap")

# NVML clocks-event (throttle) reason bits
_REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}


def _physical_index(local_index: int) -> int:
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis:
        try:
            return int(vis.split(",")[local_index])
        except (ValueError, IndexError):
            pass
    return local_index


class ClockSampler:
    def __init__(self, gpu_index: int = 0, period_ms: float = 2.0):
        self.gpu_index = _physical_index(gpu_index)
        self.period_s = period_ms / 1e3
        self._samples: list = []
        self._stop = threading.Event()
        self._thread = None
        self._proc = None
        self._path = None
        self._max = None

    # -- NVML thread ---------------------------------------------------------------------------
    def _nvml_loop(self, nv, h):
        while not self._stop.is_set():
            try:
                sm = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
                pw = nv.nvmlDeviceGetPowerUsage(h) / 1e3
                try:
                    rs = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                except Exception:
                    rs = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                self._samples.append((sm, pw, rs))
            except Exception:
                pass
            time.sleep(self.period_s)

    def start(self) -> "ClockSampler":
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.gpu_index)
            self._max = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            self._thread = threading.Thread(target=self._nvml_loop, args=(nv, h), daemon=True)
            self._thread.start()
            return self
        except Exception:
            self._thread = None
        smi = shutil.which("nvidia-smi")
        if smi:
            fd, self._path = tempfile.mkstemp(prefix="ainic_clocks_", suffix=".csv")
            os.close(fd)
            self._fh = open(self._path, "w")
            try:
                self._proc = subprocess.Popen([smi, f"--query-gpu={_QUERY}", "--format=csv,noheader,nounits", "-i",
                                               str(self.gpu_index), "-lms", "100"], stdout=self._fh,
                                              stderr=subprocess.DEVNULL)
            except Exception:
                self._proc = None
        return self

    def mark(self) -> None:
        """The timed region starts here: samples taken earlier (warm-up) are only used if the region turns
        out too short to collect two of its own."""
        self._mark = len(self._samples)

    def stop(self) -> dict:
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0, "source": None}
        sm, pw, reasons = [], [], set()
        if self._thread is not None:
            self._stop.set()
            self._thread.join(timeout=2)
            mark = getattr(self, "_mark", 0)
            use = self._samples[mark:] if len(self._samples) - mark >= 2 else self._samples
            for s, p, r in use:
                sm.append(float(s))
                pw.append(float(p))
                for bit, name in _REASONS.items():
                    if r & bit:
                        reasons.add(name)
            out["sm_max_mhz"] = float(self._max) if self._max else None
            out["source"] = "nvml"
        elif self._proc is not None:
            self._proc.terminate()      # our own child, by handle
            try:
                self._proc.wait(timeout=5)
            except Exception:
                self._proc.kill()
            self._fh.close()
            mx = []
            try:
                with open(self._path) as f:
                    for line in f:
                        p = [x.strip() for x in line.split(",")]
                        if len(p) < 9:
                            continue
                        try:
                            sm.append(float(p[1])); mx.append(float(p[2])); pw.append(float(p[3]))
                        except ValueError:
                            continue
                        for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), p[5:9]):
                            if v.lower().startswith("active"):
                                reasons.add(name)
                os.unlink(self._path)
            except OSError:
                pass
            out["sm_max_mhz"] = max(mx) if mx else None
            out["source"] = "nvidia-smi"
        if sm:
            thr = (max(pw) + min(pw)) / 2 if pw else 0      # "under load" = upper half of the power range
            loaded = [s for s, w in zip(sm, pw) if w >= thr] or sm
            out.update(sm_mhz=statistics.median(loaded), reasons=sorted(reasons), samples=len(sm),
                       power_w_max=max(pw) if pw else None)
        return out
