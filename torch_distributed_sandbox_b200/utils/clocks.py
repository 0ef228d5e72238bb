"""SM-clock / throttle sampling during a timed region (B200_PROFILING.md "clocks line")."""
from __future__ import annotations

import os
import signal
import statistics
import subprocess
import tempfile
from typing import Dict, Optional

_QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
          "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
          "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")


class ClockSampler:
    """``with ClockSampler(gpu_index) as c: ...timed region...`` then ``c.summary()``."""

    def __init__(self, gpu_index: int = 0, period_ms: int = 100):
        self.gpu_index, self.period_ms = gpu_index, period_ms
        self._proc: Optional[subprocess.Popen] = None
        self._file = None

    def __enter__(self):
        try:
            self._file = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
            self._proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={_QUERY}", "--format=csv,noheader,nounits", "-i", str(self.gpu_index),
                 "-lms", str(self.period_ms)], stdout=self._file, stderr=subprocess.DEVNULL)
        except (OSError, FileNotFoundError):
            self._proc = None
        return self

    def __exit__(self, *exc):
        if self._proc is not None:
            self._proc.send_signal(signal.SIGINT)
            try:
                self._proc.wait(timeout=3)
            except subprocess.TimeoutExpired:
                self._proc.kill()
        return False

    def summary(self) -> Dict:
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self._file is None:
            return out
        try:
            self._file.flush()
            rows = [l.strip().split(", ") for l in open(self._file.name) if l.strip()]
            os.unlink(self._file.name)
        except OSError:
            return out
        sm, names = [], ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        reasons = set()
        for r in rows:
            if len(r) < 9:
                continue
            try:
                sm.append(float(r[1]))
                out["sm_max_mhz"] = float(r[2])
                out["power_w_max"] = max(out.get("power_w_max", 0.0), float(r[3]))
            except ValueError:
                continue
            for name, v in zip(names, r[5:9]):
                if v.strip().lower() == "active":
                    reasons.add(name)
        if sm:
            # under-load samples only: drop the idle tail/head (well below the median of the top half)
            top = sorted(sm)[len(sm) // 2:]
            out["sm_mhz"] = statistics.median(top)
            out["samples"] = len(sm)
        out["reasons"] = sorted(reasons)
        return out
