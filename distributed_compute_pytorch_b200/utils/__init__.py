"""Utilities: synthetic datasets, device timers, clock sampling, tracing."""
from .data import SyntheticMNIST, SyntheticImages, SyntheticTokens  # noqa: F401
from .timing import CudaTimer, ClockSampler  # noqa: F401
from .trace import Tracer  # noqa: F401
from .checkpoint import save_model, load_model, save_training_state, load_training_state  # noqa: F401
