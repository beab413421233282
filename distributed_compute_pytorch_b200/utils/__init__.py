"""Utilities: synthetic datasets, device timers, clock sampling, checkpoint helpers."""
from .data import SyntheticMNIST, SyntheticImages, SyntheticTokens  # noqa: F401
from .timing import CudaTimer, ClockSampler  # noqa: F401
