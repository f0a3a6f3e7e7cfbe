"""Utilities: timing / clocks / tracing."""
from .timing import ClockSampler, PhaseTracer, time_steps  # noqa: F401
