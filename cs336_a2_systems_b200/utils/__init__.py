"""Timing, NVTX, clock sampling and logging helpers."""
from . import clocks, nvtx, timing

__all__ = ["clocks", "nvtx", "timing"]
