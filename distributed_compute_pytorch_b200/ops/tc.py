"""tcgen05 / TMEM / TMA tensor-core paths (csrc/ops/gemm_tc.cu, conv_tc.cu): shape gating + launchers.

``*_supported`` decide per shape whether the tensor-core kernel can take the op; everything else
falls back to the CUDA-core kernels in conv_linear_simt.cu.
"""
from __future__ import annotations

import ctypes as C

import torch

_declared = False


def declare(lib) -> None:
    global _declared
    _declared = True


def linear_supported(M: int, N: int, K: int, dtype) -> bool:
    return False


def conv_fwd_supported(p, dtype) -> bool:
    return False


def conv_dgrad_supported(p, dtype) -> bool:
    return False


def conv_wgrad_supported(p, dtype) -> bool:
    return False
