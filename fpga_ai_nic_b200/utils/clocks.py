"""Sample SM clocks and throttle reasons with nvidia-smi while a timed region runs
(the recipe's clocks line, B200_PROFILING.md)."""
from __future__ import annotations

import os
import shutil
import statistics
import subprocess
import tempfile

_QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
          "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
          "clocks_event_reasons.sw_power_cap")


class ClockSampler:
    def __init__(self, gpu_index: int = 0, period_ms: int = 100):
        self.gpu_index = gpu_index
        self.period_ms = period_ms
        self.proc = None
        self.path = None

    def start(self) -> "ClockSampler":
        smi = shutil.which("nvidia-smi")
        if not smi:
            return self
        fd, self.path = tempfile.mkstemp(prefix="ainic_clocks_", suffix=".csv")
        os.close(fd)
        self._fh = open(self.path, "w")
        try:
            self.proc = subprocess.Popen([smi, f"--query-gpu={_QUERY}", "--format=csv,noheader,nounits", "-i",
                                          str(self.gpu_index), "-lms", str(self.period_ms)], stdout=self._fh,
                                         stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None
        return self

    def stop(self) -> dict:
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        self.proc.terminate()   # our own child, by handle
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        self._fh.close()
        sm, mx, pw, reasons = [], [], [], set()
        try:
            with open(self.path) as f:
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
            os.unlink(self.path)
        except OSError:
            pass
        if sm:
            # "under load" = samples in the upper half of the power range
            thr = (max(pw) + min(pw)) / 2 if pw else 0
            loaded = [s for s, w in zip(sm, pw) if w >= thr] or sm
            out.update(sm_mhz=statistics.median(loaded), sm_max_mhz=max(mx), reasons=sorted(reasons), samples=len(sm),
                       power_w_max=max(pw) if pw else None)
        return out
