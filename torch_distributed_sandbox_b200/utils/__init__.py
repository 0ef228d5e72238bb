"""Utilities: clock sampling during timed regions (clocks.py), NVTX ranges and device step timers (trace.py)."""
from . import clocks, trace  # noqa: F401
