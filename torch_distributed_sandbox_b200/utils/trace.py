"""Tracing / profiling helpers.

The reference has no tracing at all: it prints ``datetime.now()`` deltas taken on the host without a
device synchronise (mnist_onegpu.py:61,83-84; mnist_distributed.py:83,108-109).  Here:

* ``nvtx_range(name)``  - NVTX range around a phase (forward / backward / optimizer / all-reduce launch);
  a no-op unless ``B200DP_NVTX=1`` (or ``enable_nvtx(True)``), so the hot path pays one dict lookup.
* ``StepTimer``         - CUDA-event timer on the launching stream; ``summary()`` reduces the per-step times
  over the process group with MAX, which is how every multi-GPU number in this repo is reported.
* ``phase_times(engine, ...)`` lives in tools/ddp_breakdown.py (per-kernel events).

ncu / SASS capture scripts: tools/gpu_ncu.sh, tools/ncu_summary.py (B200_PROFILING.md recipe).
"""
from __future__ import annotations

import contextlib
import os
from typing import Dict, List, Optional

import torch

_NVTX = os.environ.get("B200DP_NVTX", "0") == "1"


def enable_nvtx(on: bool = True) -> None:
    global _NVTX
    _NVTX = bool(on)


def nvtx_enabled() -> bool:
    return _NVTX


@contextlib.contextmanager
def nvtx_range(name: str):
    """NVTX range when enabled and CUDA is present; otherwise nothing."""
    if _NVTX and torch.cuda.is_available():
        torch.cuda.nvtx.range_push(name)
        try:
            yield
        finally:
            torch.cuda.nvtx.range_pop()
    else:
        yield


class StepTimer:
    """Device-side step timing.  ``with timer.step(): ...`` brackets one training step with CUDA events
    on the current stream (host ``perf_counter`` when there is no GPU, so the CPU/gloo path can use the same
    code); ``summary(group)`` returns mean/min/max ms over the recorded steps, each step taken as the MAX over
    ranks - a step is as slow as its slowest rank."""

    def __init__(self, device: Optional[torch.device] = None, skip: int = 0):
        self.cuda = torch.cuda.is_available() and (device is None or torch.device(device).type == "cuda")
        self.skip = skip
        self._events: List = []
        self._host: List[float] = []

    @contextlib.contextmanager
    def step(self):
        if self.cuda:
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            yield
            e.record()
            self._events.append((s, e))
        else:
            import time
            t0 = time.perf_counter()
            yield
            self._host.append((time.perf_counter() - t0) * 1e3)

    def times_ms(self) -> List[float]:
        if self.cuda:
            torch.cuda.synchronize()
            t = [s.elapsed_time(e) for s, e in self._events]
        else:
            t = list(self._host)
        return t[self.skip:]

    def summary(self, group=None) -> Dict[str, float]:
        t = self.times_ms()
        if not t:
            return {"steps": 0}
        if group is not None and group.size() > 1:
            from .. import distributed as dist
            buf = torch.tensor(t, dtype=torch.float64)
            if self.cuda and getattr(group.backend, "name", "") == "native":
                buf = buf.to(torch.float32).cuda()
            dist.all_reduce(buf, dist.ReduceOp.MAX, group=group)
            t = [float(x) for x in buf.cpu()]
        return {"steps": len(t), "mean_ms": sum(t) / len(t), "min_ms": min(t), "max_ms": max(t)}
