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


_NVML_REASONS = ((0x8, "hw_slowdown"), (0x40, "hw_thermal_slowdown"), (0x20, "sw_thermal_slowdown"),
                 (0x4, "sw_power_cap"))


class ClockSampler:
    """``with ClockSampler(gpu_index) as c: ...timed region...`` then ``c.summary()``.

    Samples in-process through NVML (a polling thread, ~5 ms period: a timed region of a few tens of
    milliseconds still gets samples, and eight ranks do not each fork an ``nvidia-smi``); falls back to an
    ``nvidia-smi -lms`` child when the NVML binding is missing."""

    def __init__(self, gpu_index: int = 0, period_ms: int = 100):
        self.gpu_index, self.period_ms = gpu_index, period_ms
        self._proc: Optional[subprocess.Popen] = None
        self._file = None
        self._thread = None
        self._stop = None
        self._nvml_rows = None

    def _start_nvml(self) -> bool:
        try:
            import threading
            import pynvml
            pynvml.nvmlInit()
            h = pynvml.nvmlDeviceGetHandleByIndex(self.gpu_index)
            max_sm = float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))
            reasons_fn = getattr(pynvml, "nvmlDeviceGetCurrentClocksEventReasons", None) or \
                getattr(pynvml, "nvmlDeviceGetCurrentClocksThrottleReasons")
        except Exception:
            return False
        self._nvml_rows, self._stop = [], threading.Event()
        period = min(self.period_ms, 5) / 1e3

        def poll():
            while True:
                try:
                    sm = float(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM))
                    mask = int(reasons_fn(h))
                    try:
                        pw = pynvml.nvmlDeviceGetPowerUsage(h) / 1e3
                    except Exception:
                        pw = 0.0
                    self._nvml_rows.append((sm, max_sm, pw, mask))
                except Exception:
                    pass
                if self._stop.wait(period):
                    return

        self._thread = threading.Thread(target=poll, daemon=True)
        self._thread.start()
        return True

    def __enter__(self):
        if self._start_nvml():
            return self
        try:
            self._file = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
            self._proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={_QUERY}", "--format=csv,noheader,nounits", "-i", str(self.gpu_index),
                 "-lms", str(self.period_ms)], stdout=self._file, stderr=subprocess.DEVNULL)
        except (OSError, FileNotFoundError):
            self._proc = None
        return self

    def __exit__(self, *exc):
        if self._thread is not None:
            self._stop.set()
            self._thread.join(timeout=2)
            return False
        if self._proc is not None:
            self._proc.send_signal(signal.SIGINT)
            try:
                self._proc.wait(timeout=3)
            except subprocess.TimeoutExpired:
                self._proc.kill()
        return False

    def summary(self) -> Dict:
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self._nvml_rows is not None:
            rows = list(self._nvml_rows)
            if rows:
                top = sorted(r[0] for r in rows)[len(rows) // 2:]
                out.update(sm_mhz=statistics.median(top), sm_max_mhz=rows[-1][1], samples=len(rows),
                           power_w_max=max(r[2] for r in rows), source="nvml")
                mask = 0
                for r in rows:
                    mask |= r[3]
                out["reasons"] = sorted(name for bit, name in _NVML_REASONS if mask & bit)
            return out
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
