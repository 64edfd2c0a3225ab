"""Timing, clock sampling, metrics and rank-file helpers."""
from .clocks import ClockSampler
from .metrics import gcups, roofline_fraction
from .timing import flush_l2

__all__ = ["ClockSampler", "flush_l2", "gcups", "roofline_fraction"]
