"""Timing, tracing and clock-sampling helpers."""
from .clocks import ClockSampler
from .trace import PhaseTimer

__all__ = ["ClockSampler", "PhaseTimer"]
