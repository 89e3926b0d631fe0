"""Utilities: clock/throttle sampling for benchmarks (`clocks`), and an executable model of
the ring / flag protocol for CPU-side checking (`protocol_model`)."""
from .clocks import ClockSampler  # noqa: F401
