"""Utilities: timing / clocks, input format helpers."""
from .timing import ClockSampler, time_steps  # noqa: F401
