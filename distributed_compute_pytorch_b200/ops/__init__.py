"""Operators: nn-style layers whose CUDA path is the hand-written sm_100a kernels in csrc/ops and
whose CPU path (plumbing tests only) is the plain PyTorch op."""
from .layers import (  # noqa: F401
    Conv2d, Linear, BatchNorm1d, BatchNorm2d, SyncBatchNorm, Dropout, Dropout2d, MaxPool2d, ReLU, LayerNorm, GELU,
    cross_entropy, nll_loss, log_softmax, correct_count,
)
from .optim import Adadelta, SGD, Adam, StepLR  # noqa: F401
