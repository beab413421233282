"""CUDA operator front-end: ``torch.autograd.Function`` wrappers over the hand-written sm_100a kernels
in ``libdcp_cuda.so`` (csrc/ops).  PyTorch stays the tensor/autograd layer (allocation, graph, streams);
every FLOP and byte of the listed ops runs in our kernels - the role ATen/cuDNN/cuBLAS play for the
reference model (ref:main.py:31-45,61-63).  Kernels are launched on the current torch stream through
the C ABI, so everything here can be captured in a CUDA graph.
"""
from __future__ import annotations

import ctypes as C
import os
import struct
from typing import List, Optional, Tuple

import torch
from torch.autograd import Function

from .._native import cuda_lib

_DT = {torch.float32: 0, torch.bfloat16: 1}
CL = torch.channels_last

launch_count = 0          # kernels launched through this module (bench.py reports it)
graph_step: Optional[torch.Tensor] = None   # int64 device counter bumped once per captured step (dropout streams)
_dropout_calls = 0
USE_TC = os.environ.get("DCP_DISABLE_TC") != "1"


class ConvP(C.Structure):
    _fields_ = [(n, C.c_int) for n in ("N", "H", "W", "C", "K", "R", "S", "stride", "pad", "Ho", "Wo")]


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = cuda_lib()
        vp, i, l, f, d, ull = C.c_void_p, C.c_int, C.c_long, C.c_float, C.c_double, C.c_ulonglong
        sigs = {
            "dcp_relu_fwd": [vp, vp, l, i, vp], "dcp_relu_bwd": [vp, vp, vp, l, i, vp], "dcp_add_relu_fwd": [vp, vp, vp, l, i, vp],
            "dcp_dropout": [vp, vp, l, f, ull, ull, vp, i, l, l, i, vp],
            "dcp_maxpool_fwd": [vp, vp, vp, i, i, i, i, i, i, i, i, i, i, i, vp],
            "dcp_maxpool_bwd": [vp, vp, vp, i, i, i, i, i, i, i, i, i, i, i, vp],
            "dcp_gap_fwd": [vp, vp, i, i, i, i, vp], "dcp_gap_bwd": [vp, vp, i, i, i, i, vp],
            "dcp_correct_count": [vp, vp, vp, i, i, i, vp],
            "dcp_bn_stats": [vp, vp, l, i, i, vp],
            "dcp_bn_apply": [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, l, i, d, f, f, i, i, i, vp],
            "dcp_bn_bwd_reduce": [vp, vp, vp, vp, vp, vp, l, i, i, i, vp],
            "dcp_bn_bwd_apply": [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, l, i, d, i, i, vp],
            "dcp_log_softmax_fwd": [vp, vp, i, i, i, vp], "dcp_log_softmax_bwd": [vp, vp, vp, i, i, i, vp],
            "dcp_nll_fwd": [vp, vp, vp, i, i, f, i, vp], "dcp_nll_bwd": [vp, vp, vp, i, i, f, i, vp],
            "dcp_ce_fwd": [vp, vp, vp, vp, i, i, f, i, vp], "dcp_ce_bwd": [vp, vp, vp, vp, vp, i, i, f, i, vp],
            "dcp_optim_advance": [vp, vp], "dcp_optim_step": [vp, i, C.c_longlong, vp, i, i, vp],
            "dcp_conv2d_fwd_simt": [vp, vp, vp, vp, ConvP, i, i, vp], "dcp_conv2d_dgrad_simt": [vp, vp, vp, ConvP, i, vp],
            "dcp_conv2d_wgrad_simt": [vp, vp, vp, vp, ConvP, i, vp],
            "dcp_linear_fwd_simt": [vp, vp, vp, vp, i, i, i, i, i, vp], "dcp_linear_dgrad_simt": [vp, vp, vp, i, i, i, i, vp],
            "dcp_linear_wgrad_simt": [vp, vp, vp, vp, i, i, i, i, vp],
            "dcp_colsum": [vp, vp, l, i, i, vp], "dcp_cast_from_f32": [vp, vp, l, i, vp],
            "dcp_counter_advance": [vp, vp],
            "dcp_layernorm_fwd": [vp, vp, vp, vp, vp, vp, i, i, f, i, vp],
            "dcp_layernorm_bwd": [vp, vp, vp, vp, vp, vp, vp, vp, i, i, i, vp],
            "dcp_gelu_fwd": [vp, vp, l, i, vp], "dcp_gelu_bwd": [vp, vp, vp, l, i, vp],
            "dcp_conv_flat_tc": [vp, vp, vp, vp, vp, l, i, i, i, vp, i, i, i, i, i, i, vp],
            "dcp_add_layernorm_fwd": [vp, vp, vp, vp, vp, vp, vp, vp, i, i, C.c_float, i, vp],
            "dcp_add_layernorm_bwd": [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i, i, i, vp],
            "dcp_gelu_bwd_colsum": [vp, vp, vp, vp, l, i, i, vp],
            "dcp_embed_fwd": [vp, vp, vp, vp, l, i, i, i, vp], "dcp_embed_bwd": [vp, vp, vp, vp, l, i, i, i, vp],
            "dcp_attn_fwd": [vp, vp, vp, i, i, i, vp],
            "dcp_attn_bwd": [vp, vp, vp, vp, vp, vp, vp, i, i, i, vp],
            "dcp_conv_flat_pool_tc": [vp, vp, vp, vp, vp, l, i, i, i, vp, i, i, i, i, f, ull, ull, vp, i, i, vp],
            "dcp_conv_wgrad_flat_tc": [vp, vp, vp, l, i, i, i, vp],
            "dcp_convnet_conv1_fwd": [vp, vp, vp, vp, i, vp],
            "dcp_convnet_pool_fwd": [vp, vp, vp, i, f, ull, ull, vp, i, vp],
            "dcp_convnet_pool_bwd": [vp, vp, vp, vp, i, f, ull, ull, vp, i, vp],
            "dcp_convnet_conv1_wgrad": [vp, vp, vp, vp, i, vp],
            "dcp_convnet_conv1_wgrad_allreduce": [vp, vp, vp, vp, vp, i, vp, i, i, vp],
            "dcp_convnet_head": [vp] * 17 + [i, f, f, f, ull, ull, vp, i, vp],
            "dcp_convnet_head_dp": [vp] * 16 + [i, f, f, f, ull, ull, vp, i, vp],
            "dcp_convnet_conv1_wgrad_push": [vp, vp, vp, i, vp, C.c_uint, C.c_uint, i, i, vp, vp, vp],
            "dcp_convnet_conv1_wgrad_ll": [vp, vp, vp, i, vp, vp, vp, i, vp, vp, vp],
            "dcp_splitk_reduce_push": [vp, vp, i, l, vp, C.c_uint, C.c_uint, i, i, vp, vp, i, vp],
            "dcp_dp_tail": [vp, vp, i, C.c_longlong, vp, i, vp, vp, C.c_uint, i, vp, i, vp, vp, i, vp],
        }
        for name, args in sigs.items():
            fn = getattr(_lib, name)
            fn.restype, fn.argtypes = C.c_int, args
        from . import tc
        tc.declare(_lib)
    return _lib


def advance_graph_step() -> None:
    """Bump the device-side step counter (first kernel of a captured training step)."""
    if graph_step is not None:
        _chk(lib().dcp_counter_advance(graph_step.data_ptr(), _stream()))


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _chk(rc: int) -> None:
    global launch_count
    launch_count += 1
    if rc != 0:
        raise RuntimeError(f"libdcp_cuda: {(lib().dcp_last_error() or b'').decode()}")


def _dt(t: torch.Tensor) -> int:
    try:
        return _DT[t.dtype]
    except KeyError:
        raise TypeError(f"dtype {t.dtype} unsupported by the sm_100a kernels (float32 / bfloat16)")


def _p(t: Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


def _nhwc(x: torch.Tensor) -> torch.Tensor:
    return x if x.is_contiguous(memory_format=CL) else x.contiguous(memory_format=CL)


# ------------------------------------------------------------------------------------ elementwise --
class _ReLU(Function):
    @staticmethod
    def forward(ctx, x):
        x = x if (x.is_contiguous() or x.is_contiguous(memory_format=CL)) else x.contiguous()
        y = torch.empty_like(x)
        _chk(lib().dcp_relu_fwd(x.data_ptr(), y.data_ptr(), x.numel(), _dt(x), _stream()))
        ctx.save_for_backward(y)
        return y

    @staticmethod
    def backward(ctx, dy):
        (y,) = ctx.saved_tensors
        dy = _like_layout(dy, y)
        dx = torch.empty_like(y)
        _chk(lib().dcp_relu_bwd(y.data_ptr(), dy.data_ptr(), dx.data_ptr(), y.numel(), _dt(y), _stream()))
        return dx


def _like_layout(t: torch.Tensor, ref: torch.Tensor) -> torch.Tensor:
    """Make ``t`` dense with the same physical layout as ``ref`` (so flat indexing agrees)."""
    if t.stride() == ref.stride():
        return t
    out = torch.empty_like(ref)
    out.copy_(t)
    return out


def relu(x):
    return _ReLU.apply(x)


class _AddReLU(Function):
    @staticmethod
    def forward(ctx, a, b):
        b = _like_layout(b, a)
        y = torch.empty_like(a)
        _chk(lib().dcp_add_relu_fwd(a.data_ptr(), b.data_ptr(), y.data_ptr(), a.numel(), _dt(a), _stream()))
        ctx.save_for_backward(y)
        return y

    @staticmethod
    def backward(ctx, dy):
        (y,) = ctx.saved_tensors
        dy = _like_layout(dy, y)
        dx = torch.empty_like(y)
        _chk(lib().dcp_relu_bwd(y.data_ptr(), dy.data_ptr(), dx.data_ptr(), y.numel(), _dt(y), _stream()))
        return dx, dx


def add_relu(a, b):
    a = a if (a.is_contiguous() or a.is_contiguous(memory_format=CL)) else a.contiguous()
    return _AddReLU.apply(a, b)


class _Dropout(Function):
    @staticmethod
    def forward(ctx, x, p, mode, inner, hw, seed, stream_id):
        y = torch.empty_like(x)
        ctx.cfg = (p, mode, inner, hw, seed, stream_id, x.stride())
        _chk(lib().dcp_dropout(x.data_ptr(), y.data_ptr(), x.numel(), p, seed, stream_id, _p(graph_step), mode, inner, hw, _dt(x), _stream()))
        return y

    @staticmethod
    def backward(ctx, dy):
        p, mode, inner, hw, seed, stream_id, stride = ctx.cfg
        if dy.stride() != stride:  # the mask is indexed by physical position: match the forward layout
            d2 = torch.empty_strided(dy.shape, stride, dtype=dy.dtype, device=dy.device)
            d2.copy_(dy)
            dy = d2
        dx = torch.empty_like(dy)
        _chk(lib().dcp_dropout(dy.data_ptr(), dx.data_ptr(), dy.numel(), p, seed, stream_id, _p(graph_step), mode, inner, hw, _dt(dy), _stream()))
        return dx, None, None, None, None, None, None


def dropout(x, p: float, training: bool, per_channel: bool = False):
    """Philox dropout; ``per_channel`` drops whole (n, c) planes of a 4-D tensor (Dropout2d)."""
    global _dropout_calls
    if not training or p == 0.0:
        return x
    if p >= 1.0:
        return x * 0
    x = x if (x.is_contiguous() or (x.dim() == 4 and x.is_contiguous(memory_format=CL))) else x.contiguous()
    mode, inner, hw = 0, 1, 1
    if per_channel and x.dim() == 4:
        _, Cc, H, W = x.shape
        mode, inner, hw = (2, Cc, H * W) if x.is_contiguous() else (1, Cc, H * W)
    _dropout_calls += 1
    seed = torch.initial_seed() & 0xFFFFFFFFFFFFFFFF
    return _Dropout.apply(x, float(p), mode, inner, hw, seed, _dropout_calls)


class _MaxPool(Function):
    @staticmethod
    def forward(ctx, x, k, s, pad, out_nchw):
        x = _nhwc(x)
        N, Cc, H, W = x.shape
        Ho, Wo = (H + 2 * pad - k) // s + 1, (W + 2 * pad - k) // s + 1
        y = torch.empty((N, Cc, Ho, Wo), dtype=x.dtype, device=x.device, memory_format=torch.contiguous_format if out_nchw else CL)
        arg = torch.empty(N * Ho * Wo * Cc, dtype=torch.uint8, device=x.device)
        _chk(lib().dcp_maxpool_fwd(x.data_ptr(), y.data_ptr(), arg.data_ptr(), N, H, W, Cc, Ho, Wo, k, s, pad, int(out_nchw), _dt(x), _stream()))
        ctx.save_for_backward(arg)
        ctx.cfg = (N, H, W, Cc, Ho, Wo, k, s, pad, out_nchw)
        return y

    @staticmethod
    def backward(ctx, dy):
        (arg,) = ctx.saved_tensors
        N, H, W, Cc, Ho, Wo, k, s, pad, out_nchw = ctx.cfg
        dy = dy.contiguous() if out_nchw else _nhwc(dy)
        dx = torch.empty((N, Cc, H, W), dtype=dy.dtype, device=dy.device, memory_format=CL)
        _chk(lib().dcp_maxpool_bwd(dy.data_ptr(), arg.data_ptr(), dx.data_ptr(), N, H, W, Cc, Ho, Wo, k, s, pad, int(out_nchw), _dt(dy), _stream()))
        return dx, None, None, None, None


def max_pool2d(x, k: int, s: int, pad: int = 0, out_nchw: bool = False):
    return _MaxPool.apply(x, k, s, pad, out_nchw)


class _GAP(Function):
    @staticmethod
    def forward(ctx, x):
        x = _nhwc(x)
        N, Cc, H, W = x.shape
        y = torch.empty((N, Cc), dtype=x.dtype, device=x.device)
        _chk(lib().dcp_gap_fwd(x.data_ptr(), y.data_ptr(), N, H * W, Cc, _dt(x), _stream()))
        ctx.cfg = (N, Cc, H, W)
        return y

    @staticmethod
    def backward(ctx, dy):
        N, Cc, H, W = ctx.cfg
        dy = dy.contiguous()
        dx = torch.empty((N, Cc, H, W), dtype=dy.dtype, device=dy.device, memory_format=CL)
        _chk(lib().dcp_gap_bwd(dy.data_ptr(), dx.data_ptr(), N, H * W, Cc, _dt(dy), _stream()))
        return dx


def global_avg_pool(x):
    return _GAP.apply(x)


def correct_count(logits: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
    x = logits.detach().contiguous()
    out = torch.empty((), dtype=torch.int64, device=x.device)
    _chk(lib().dcp_correct_count(x.data_ptr(), target.contiguous().data_ptr(), out.data_ptr(), x.shape[0], x.shape[1], _dt(x), _stream()))
    return out


# -------------------------------------------------------------------------------------- batch norm --
class _BatchNorm(Function):
    @staticmethod
    def forward(ctx, x, gamma, beta, run_mean, run_var, nbt, training, momentum, eps, relu, sync):
        four_d = x.dim() == 4
        x = _nhwc(x) if four_d else x.contiguous()
        Cc = x.shape[1]
        M = x.numel() // Cc
        y = torch.empty_like(x)
        dev = x.device
        scratch = torch.empty(4 * Cc + 2, dtype=torch.float64, device=dev) if training else None
        mean = torch.empty(Cc, dtype=torch.float32, device=dev) if training else None
        rstd = torch.empty(Cc, dtype=torch.float32, device=dev) if training else None
        count = float(M)
        if training:
            _chk(lib().dcp_bn_stats(x.data_ptr(), scratch.data_ptr(), M, Cc, _dt(x), _stream()))
            if sync:
                count = _sync_stats(scratch, 2 * Cc, M)
        mom = 0.1 if momentum is None else float(momentum)
        _chk(lib().dcp_bn_apply(x.data_ptr(), y.data_ptr(), _p(scratch), _p(gamma), _p(beta), _p(run_mean), _p(run_var), _p(nbt),
                                _p(mean), _p(rstd), M, Cc, count, float(eps), mom, int(relu), int(training), _dt(x), _stream()))
        if training:
            ctx.save_for_backward(x, y if relu else x, gamma, mean, rstd)
            ctx.cfg = (M, Cc, relu, sync, count)
        ctx.training = training
        if not training:
            ctx.save_for_backward(x, gamma, run_mean, run_var)
            ctx.cfg = (M, Cc, relu, eps, y if relu else None)
        return y

    @staticmethod
    def backward(ctx, dy):
        if not ctx.training:
            x, gamma, run_mean, run_var = ctx.saved_tensors
            M, Cc, relu, eps, y = ctx.cfg
            shape = [1, -1] + [1] * (x.dim() - 2)
            g = dy * (y > 0) if relu else dy
            k = torch.rsqrt(run_var + eps) * (gamma if gamma is not None else 1.0)
            xhat = (x - run_mean.view(shape)) * torch.rsqrt(run_var + eps).view(shape)
            dims = [d for d in range(x.dim()) if d != 1]
            return (g * k.view(shape), (g * xhat).sum(dims) if gamma is not None else None, g.sum(dims) if gamma is not None else None) + (None,) * 8
        x, y, gamma, mean, rstd = ctx.saved_tensors
        M, Cc, relu, sync, count = ctx.cfg
        dy = _like_layout(dy, x)
        dx = torch.empty_like(x)
        dev = x.device
        scratch = torch.empty(4 * Cc + 2, dtype=torch.float64, device=dev)
        dgamma = torch.empty(Cc, dtype=torch.float32, device=dev) if gamma is not None else None
        dbeta = torch.empty(Cc, dtype=torch.float32, device=dev) if gamma is not None else None
        _chk(lib().dcp_bn_bwd_reduce(x.data_ptr(), y.data_ptr(), dy.data_ptr(), mean.data_ptr(), rstd.data_ptr(), scratch.data_ptr(), M, Cc,
                                     int(relu), _dt(x), _stream()))
        if sync:
            _sync_stats(scratch, 2 * Cc, None)
        _chk(lib().dcp_bn_bwd_apply(x.data_ptr(), y.data_ptr(), dy.data_ptr(), dx.data_ptr(), mean.data_ptr(), rstd.data_ptr(), _p(gamma),
                                    scratch.data_ptr(), _p(dgamma), _p(dbeta), M, Cc, count, int(relu), _dt(x), _stream()))
        return (dx, dgamma, dbeta) + (None,) * 8


def _sync_stats(scratch: torch.Tensor, n: int, local_rows: Optional[int]) -> float:
    """SyncBatchNorm: sum the first ``n`` doubles of ``scratch`` (and the row count) across ranks with
    the one-shot P2P all_reduce on its own signal channel.  Returns the global row count."""
    from .. import distributed as dist
    from ..parallel.p2p_backend import CH_BN
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return float(local_rows or 0)
    be = dist._world().cuda
    be.all_reduce(scratch[:n], dist.ReduceOp.SUM, algo=1, chan=CH_BN)
    if local_rows is None:
        return 0.0
    return float(local_rows * dist.get_world_size())  # equal per-rank batches (DistributedSampler pads to equal shards)


def batch_norm(x, gamma, beta, run_mean, run_var, nbt, training, momentum, eps, relu, sync=False):
    return _BatchNorm.apply(x, gamma, beta, run_mean, run_var, nbt, training, momentum, eps, relu, sync)


# ------------------------------------------------------------------------------------------- loss --
class _LogSoftmax(Function):
    @staticmethod
    def forward(ctx, x):
        x = x.contiguous()
        y = torch.empty_like(x)
        _chk(lib().dcp_log_softmax_fwd(x.data_ptr(), y.data_ptr(), x.shape[0], x.shape[1], _dt(x), _stream()))
        ctx.save_for_backward(y)
        return y

    @staticmethod
    def backward(ctx, dy):
        (y,) = ctx.saved_tensors
        dy = dy.contiguous()
        dx = torch.empty_like(y)
        _chk(lib().dcp_log_softmax_bwd(y.data_ptr(), dy.data_ptr(), dx.data_ptr(), y.shape[0], y.shape[1], _dt(y), _stream()))
        return dx


def log_softmax(x):
    return _LogSoftmax.apply(x)


class _NLL(Function):
    @staticmethod
    def forward(ctx, logp, target, mean):
        logp = logp.contiguous()
        target = target.contiguous()
        M, Cc = logp.shape
        loss = torch.empty((), dtype=torch.float32, device=logp.device)
        scale = 1.0 / M if mean else 1.0
        _chk(lib().dcp_nll_fwd(logp.data_ptr(), target.data_ptr(), loss.data_ptr(), M, Cc, scale, _dt(logp), _stream()))
        ctx.save_for_backward(target)
        ctx.cfg = (M, Cc, scale, logp.dtype)
        return loss

    @staticmethod
    def backward(ctx, g):
        (target,) = ctx.saved_tensors
        M, Cc, scale, dtype = ctx.cfg
        g = g.contiguous().float()
        d = torch.empty((M, Cc), dtype=dtype, device=g.device)
        _chk(lib().dcp_nll_bwd(target.data_ptr(), g.data_ptr(), d.data_ptr(), M, Cc, scale, _DT[dtype], _stream()))
        return d, None, None


def nll_loss(logp, target, reduction: str = "mean"):
    if reduction not in ("mean", "sum"):
        raise ValueError("nll_loss: reduction must be 'mean' or 'sum'")
    return _NLL.apply(logp, target, reduction == "mean")


class _CrossEntropy(Function):
    @staticmethod
    def forward(ctx, x, target, mean):
        x = x.contiguous()
        target = target.contiguous()
        M, Cc = x.shape
        loss = torch.empty((), dtype=torch.float32, device=x.device)
        lse = torch.empty(M, dtype=torch.float32, device=x.device)
        scale = 1.0 / M if mean else 1.0
        _chk(lib().dcp_ce_fwd(x.data_ptr(), target.data_ptr(), loss.data_ptr(), lse.data_ptr(), M, Cc, scale, _dt(x), _stream()))
        ctx.save_for_backward(x, target, lse)
        ctx.scale = scale
        return loss

    @staticmethod
    def backward(ctx, g):
        x, target, lse = ctx.saved_tensors
        g = g.contiguous().float()
        dx = torch.empty_like(x)
        _chk(lib().dcp_ce_bwd(x.data_ptr(), target.data_ptr(), lse.data_ptr(), g.data_ptr(), dx.data_ptr(), x.shape[0], x.shape[1], ctx.scale,
                              _dt(x), _stream()))
        return dx, None, None


def cross_entropy(x, target, reduction: str = "mean"):
    return _CrossEntropy.apply(x, target, reduction == "mean")


# ---------------------------------------------------------------------------------- conv / linear --
def _convp(x_shape, w_shape, stride, pad) -> ConvP:
    N, Cc, H, W = x_shape
    K, _, R, S = w_shape
    Ho, Wo = (H + 2 * pad - R) // stride + 1, (W + 2 * pad - S) // stride + 1
    return ConvP(N, H, W, Cc, K, R, S, stride, pad, Ho, Wo)


class _Conv2d(Function):
    @staticmethod
    def forward(ctx, x, w, b, stride, pad, relu):
        from . import tc
        x = _nhwc(x)
        w = _nhwc(w)
        ctx.w_dtype = w.dtype
        if w.dtype != x.dtype:      # fp32 master weights, bf16 compute (autocast-style): cast per call, grads come back fp32
            w = _cast(w, x.dtype) if w.dtype == torch.float32 else w.to(x.dtype)
            b = None if b is None else (_cast(b, x.dtype) if b.dtype == torch.float32 else b.to(x.dtype))
        p = _convp(x.shape, w.shape, stride, pad)
        y = torch.empty((p.N, p.K, p.Ho, p.Wo), dtype=x.dtype, device=x.device, memory_format=CL)
        if USE_TC and tc.conv_fwd_supported(p, x.dtype):
            tc.conv2d_fwd(x, w, b, y, p, relu)
        else:
            _chk(lib().dcp_conv2d_fwd_simt(x.data_ptr(), w.data_ptr(), _p(b), y.data_ptr(), p, int(relu), _dt(x), _stream()))
        ctx.save_for_backward(x, w, y if relu else None)
        ctx.cfg = (p, relu, b is not None)
        return y

    @staticmethod
    def backward(ctx, dy):
        from . import tc
        x, w, y = ctx.saved_tensors
        p, relu, has_b = ctx.cfg
        dy = _nhwc(dy)
        if relu:
            g = torch.empty_like(dy)
            _chk(lib().dcp_relu_bwd(y.data_ptr(), dy.data_ptr(), g.data_ptr(), y.numel(), _dt(y), _stream()))
            dy = g
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty_like(x)
            if USE_TC and tc.conv_dgrad_supported(p, x.dtype):
                tc.conv2d_dgrad(dy, w, dx, p)
            else:
                _chk(lib().dcp_conv2d_dgrad_simt(dy.data_ptr(), w.data_ptr(), dx.data_ptr(), p, _dt(x), _stream()))
        if ctx.needs_input_grad[1]:
            dw32 = torch.empty_like(w, dtype=torch.float32)
            db32 = torch.empty(p.K, dtype=torch.float32, device=x.device) if has_b else None
            if USE_TC and tc.conv_wgrad_supported(p, x.dtype):
                tc.conv2d_wgrad(dy, x, dw32, db32, p)
            else:
                _chk(lib().dcp_conv2d_wgrad_simt(dy.data_ptr(), x.data_ptr(), dw32.data_ptr(), _p(db32), p, _dt(x), _stream()))
            dw = dw32 if ctx.w_dtype == torch.float32 else _cast(dw32, ctx.w_dtype)
            if has_b:
                db = db32 if ctx.w_dtype == torch.float32 else _cast(db32, ctx.w_dtype)
        return dx, dw, db, None, None, None


# Column sums that a backward kernel produced as a by-product of writing a gradient tensor: [(tensor, colsum)].  The Linear whose
# output gradient IS that tensor takes its bias gradient from here instead of reading dy once more (GELU backward -> mlp_fc,
# add+LayerNorm backward -> attn_proj / mlp_proj).  The entry keeps the tensor alive, so its address cannot be reused meanwhile.
_COLSUMS = []


def _offer_colsum(t: torch.Tensor, colsum: torch.Tensor) -> None:
    _COLSUMS.append((t, colsum))
    if len(_COLSUMS) > 4:
        del _COLSUMS[0]


def _take_colsum(dy: torch.Tensor, n: int):
    for k, (t, cs) in enumerate(_COLSUMS):
        if t.data_ptr() == dy.data_ptr() and t.numel() == dy.numel() and cs.numel() == n and t.dtype == dy.dtype:
            del _COLSUMS[k]
            return cs
    return None


# Low-precision shadows of fp32 master weights: (data_ptr, dtype) -> tensor of that dtype which its owner keeps current (the
# fused sharded optimizer re-casts a whole bucket with ONE launch after each update, parallel/zero.py).  Linear layers use the
# shadow instead of casting their weight on every forward.
_SHADOWS = {}


def register_shadow(master: torch.Tensor, shadow: torch.Tensor) -> None:
    _SHADOWS[(master.data_ptr(), shadow.dtype)] = (shadow, tuple(master.shape))


def drop_shadows(masters) -> None:
    for m in masters:
        for k in [k for k in _SHADOWS if k[0] == m.data_ptr()]:
            del _SHADOWS[k]


def _lowp(w32: torch.Tensor, dtype) -> torch.Tensor:
    ent = _SHADOWS.get((w32.data_ptr(), dtype))
    if ent is not None and ent[1] == tuple(w32.shape):
        return ent[0]
    return _cast(w32, dtype)


def _cast(t32: torch.Tensor, dtype) -> torch.Tensor:
    out = torch.empty_like(t32, dtype=dtype)
    _chk(lib().dcp_cast_from_f32(t32.data_ptr(), out.data_ptr(), t32.numel(), _DT[dtype], _stream()))
    return out


def conv2d(x, w, b, stride, padding, relu: bool = False):
    s = stride[0] if isinstance(stride, (tuple, list)) else stride
    pd = padding[0] if isinstance(padding, (tuple, list)) else padding
    return _Conv2d.apply(x, w, b, int(s), int(pd), bool(relu))


class _Linear(Function):
    @staticmethod
    def forward(ctx, x, w, b, relu):
        from . import tc
        shape = x.shape
        x2 = x.reshape(-1, shape[-1])
        x2 = x2 if x2.is_contiguous() else x2.contiguous()
        w = w if w.is_contiguous() else w.contiguous()
        ctx.w_dtype = w.dtype
        if w.dtype != x2.dtype:     # fp32 master weights, bf16 compute
            w = _lowp(w, x2.dtype) if w.dtype == torch.float32 else w.to(x2.dtype)
            b = None if b is None else (_lowp(b, x2.dtype) if b.dtype == torch.float32 else b.to(x2.dtype))
        M, K = x2.shape
        N = w.shape[0]
        y = torch.empty((M, N), dtype=x.dtype, device=x.device)
        if USE_TC and tc.linear_supported(M, N, K, x.dtype):
            tc.linear_fwd(x2, w, b, y, relu)
        else:
            _chk(lib().dcp_linear_fwd_simt(x2.data_ptr(), w.data_ptr(), _p(b), y.data_ptr(), M, N, K, int(relu), _dt(x2), _stream()))
        ctx.save_for_backward(x2, w, y if relu else None)
        ctx.cfg = (shape, relu, b is not None)
        return y.view(*shape[:-1], N)

    @staticmethod
    def backward(ctx, dy):
        from . import tc
        x2, w, y = ctx.saved_tensors
        shape, relu, has_b = ctx.cfg
        M, K = x2.shape
        N = w.shape[0]
        dy = dy.reshape(M, N)
        dy = dy if dy.is_contiguous() else dy.contiguous()
        if relu:
            g = torch.empty_like(dy)
            _chk(lib().dcp_relu_bwd(y.data_ptr(), dy.data_ptr(), g.data_ptr(), y.numel(), _dt(y), _stream()))
            dy = g
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty_like(x2)
            if USE_TC and tc.linear_supported(M, K, N, x2.dtype):
                tc.linear_dgrad(dy, w, dx)
            else:
                _chk(lib().dcp_linear_dgrad_simt(dy.data_ptr(), w.data_ptr(), dx.data_ptr(), M, N, K, _dt(x2), _stream()))
            dx = dx.view(shape)
        if ctx.needs_input_grad[1]:
            dw32 = torch.empty((N, K), dtype=torch.float32, device=x2.device)
            ready = _take_colsum(dy, N) if (has_b and not relu) else None      # bias gradient already summed by dy's producer
            db32 = ready if ready is not None else (torch.empty(N, dtype=torch.float32, device=x2.device) if has_b else None)
            if USE_TC and tc.linear_supported(N, K, M, x2.dtype):
                tc.linear_wgrad(dy, x2, dw32, None if ready is not None else db32)
            else:
                _chk(lib().dcp_linear_wgrad_simt(dy.data_ptr(), x2.data_ptr(), dw32.data_ptr(), None if ready is not None else _p(db32), M, N, K,
                                                 _dt(x2), _stream()))
            dw = dw32 if ctx.w_dtype == torch.float32 else _cast(dw32, ctx.w_dtype)
            if has_b:
                db = db32 if ctx.w_dtype == torch.float32 else _cast(db32, ctx.w_dtype)
        return dx, dw, db, None


def linear(x, w, b, relu: bool = False):
    return _Linear.apply(x, w, b, bool(relu))


# ------------------------------------------------------------------------------------ transformer --
class _LayerNorm(Function):
    @staticmethod
    def forward(ctx, x, gamma, beta, eps):
        shape = x.shape
        x2 = x.reshape(-1, shape[-1])
        x2 = x2 if x2.is_contiguous() else x2.contiguous()
        rows, Cc = x2.shape
        y = torch.empty_like(x2)
        mean = torch.empty(rows, dtype=torch.float32, device=x.device)
        rstd = torch.empty(rows, dtype=torch.float32, device=x.device)
        _chk(lib().dcp_layernorm_fwd(x2.data_ptr(), y.data_ptr(), gamma.data_ptr(), beta.data_ptr(), mean.data_ptr(), rstd.data_ptr(), rows, Cc,
                                     float(eps), _dt(x2), _stream()))
        ctx.save_for_backward(x2, gamma, mean, rstd)
        ctx.shape = shape
        return y.view(shape)

    @staticmethod
    def backward(ctx, dy):
        x2, gamma, mean, rstd = ctx.saved_tensors
        rows, Cc = x2.shape
        dy = dy.reshape(rows, Cc)
        dy = dy if dy.is_contiguous() else dy.contiguous()
        dx = torch.empty_like(x2)
        dg = torch.empty(Cc, dtype=torch.float32, device=x2.device)
        db = torch.empty(Cc, dtype=torch.float32, device=x2.device)
        _chk(lib().dcp_layernorm_bwd(x2.data_ptr(), dy.data_ptr(), gamma.data_ptr(), mean.data_ptr(), rstd.data_ptr(), dx.data_ptr(), dg.data_ptr(),
                                     db.data_ptr(), rows, Cc, _dt(x2), _stream()))
        return dx.view(ctx.shape), dg, db, None


def layer_norm(x, gamma, beta, eps: float = 1e-5):
    return _LayerNorm.apply(x, gamma, beta, eps)


class _AddLayerNorm(Function):
    """(x, res) -> (s, y) with s = x + res and y = LayerNorm(s): the residual add of a pre-LN transformer block fused into the
    LayerNorm that follows it (one pass over the row instead of an add kernel plus a LayerNorm).  Backward folds the gradient
    that reaches ``s`` from the residual path into the LayerNorm backward, so autograd's accumulation add disappears as well;
    x and res receive the same gradient tensor."""

    @staticmethod
    def forward(ctx, x, res, gamma, beta, eps):
        shape = x.shape
        x2 = x.reshape(-1, shape[-1]); x2 = x2 if x2.is_contiguous() else x2.contiguous()
        r2 = res.reshape(-1, shape[-1]); r2 = r2 if r2.is_contiguous() else r2.contiguous()
        rows, Cc = x2.shape
        ssum, y = torch.empty_like(x2), torch.empty_like(x2)
        mean = torch.empty(rows, dtype=torch.float32, device=x.device)
        rstd = torch.empty(rows, dtype=torch.float32, device=x.device)
        _chk(lib().dcp_add_layernorm_fwd(x2.data_ptr(), r2.data_ptr(), ssum.data_ptr(), y.data_ptr(), gamma.data_ptr(), beta.data_ptr(), mean.data_ptr(),
                                         rstd.data_ptr(), rows, Cc, float(eps), _dt(x2), _stream()))
        ctx.save_for_backward(ssum, gamma, mean, rstd)
        ctx.shape = shape
        ctx.set_materialize_grads(False)
        return ssum.view(shape), y.view(shape)

    @staticmethod
    def backward(ctx, dsum, dy):
        ssum, gamma, mean, rstd = ctx.saved_tensors
        rows, Cc = ssum.shape
        if dy is None:                                   # only the sum was used
            return dsum, dsum, None, None, None
        dy = dy.reshape(rows, Cc); dy = dy if dy.is_contiguous() else dy.contiguous()
        dx = torch.empty_like(ssum)
        dg = torch.empty(Cc, dtype=torch.float32, device=ssum.device)
        db = torch.empty(Cc, dtype=torch.float32, device=ssum.device)
        if dsum is not None:
            dsum = dsum.reshape(rows, Cc); dsum = dsum if dsum.is_contiguous() else dsum.contiguous()
        cs = torch.empty(Cc, dtype=torch.float32, device=ssum.device)
        _chk(lib().dcp_add_layernorm_bwd(ssum.data_ptr(), dy.data_ptr(), _p(dsum), gamma.data_ptr(), mean.data_ptr(), rstd.data_ptr(),
                                         dx.data_ptr(), dg.data_ptr(), db.data_ptr(), cs.data_ptr(), rows, Cc, _dt(ssum), _stream()))
        _offer_colsum(dx, cs)        # dx is the output gradient of the Linear that produced `res`: its bias gradient is ready
        dx = dx.view(ctx.shape)
        return dx, dx, dg, db, None


def add_layer_norm_supported(x: torch.Tensor) -> bool:
    vn = 4 if x.dtype == torch.float32 else 8
    return x.is_cuda and x.dtype in (torch.float32, torch.bfloat16) and x.shape[-1] % vn == 0 and x.shape[-1] // vn <= 128


def add_layer_norm(x, res, gamma, beta, eps: float = 1e-5):
    """Returns (x + res, LayerNorm(x + res))."""
    return _AddLayerNorm.apply(x, res, gamma, beta, eps)


class _Embed(Function):
    """Token + position embedding: out[b, t, :] = dtype(wte[idx[b, t]] + wpe[t]) in one gather kernel; backward scatters into
    the (dense, fp32) table gradients with vector reductions instead of the sort-based library path."""

    @staticmethod
    def forward(ctx, idx, wte, wpe, dtype):
        B, T = idx.shape
        Cc = wte.shape[1]
        idx = idx if idx.is_contiguous() else idx.contiguous()
        out = torch.empty(B, T, Cc, dtype=dtype, device=wte.device)
        _chk(lib().dcp_embed_fwd(idx.data_ptr(), wte.data_ptr(), wpe.data_ptr(), out.data_ptr(), B * T, T, Cc, _DT[dtype], _stream()))
        ctx.save_for_backward(idx)
        ctx.shapes = (wte.shape, wpe.shape)
        return out

    @staticmethod
    def backward(ctx, dy):
        (idx,) = ctx.saved_tensors
        B, T = idx.shape
        ws, ps = ctx.shapes
        dy = dy if dy.is_contiguous() else dy.contiguous()
        dwte = torch.zeros(ws, dtype=torch.float32, device=dy.device)
        dwpe = torch.zeros(ps, dtype=torch.float32, device=dy.device) if ps[0] > T else torch.empty(ps, dtype=torch.float32, device=dy.device)
        _chk(lib().dcp_embed_bwd(idx.data_ptr(), dy.data_ptr(), dwte.data_ptr(), dwpe.data_ptr(), B * T, T, ws[1], _dt(dy), _stream()))
        return None, dwte, dwpe, None


def embed(idx, wte, wpe, dtype):
    return _Embed.apply(idx, wte, wpe, dtype)


class _GELU(Function):
    @staticmethod
    def forward(ctx, x):
        x = x if x.is_contiguous() else x.contiguous()
        y = torch.empty_like(x)
        _chk(lib().dcp_gelu_fwd(x.data_ptr(), y.data_ptr(), x.numel(), _dt(x), _stream()))
        ctx.save_for_backward(x)
        return y

    @staticmethod
    def backward(ctx, dy):
        (x,) = ctx.saved_tensors
        dy = dy if dy.is_contiguous() else dy.contiguous()
        dx = torch.empty_like(x)
        Cc = x.shape[-1]
        if x.dim() >= 2 and x.dtype in (torch.float32, torch.bfloat16) and Cc % (4 if x.dtype == torch.float32 else 8) == 0 and x.numel() >= (1 << 16):
            # also sums dx over the rows: the bias gradient of the Linear that produced x
            cs = torch.empty(Cc, dtype=torch.float32, device=x.device)
            _chk(lib().dcp_gelu_bwd_colsum(x.data_ptr(), dy.data_ptr(), dx.data_ptr(), cs.data_ptr(), x.numel() // Cc, Cc, _dt(x), _stream()))
            _offer_colsum(dx, cs)
        else:
            _chk(lib().dcp_gelu_bwd(x.data_ptr(), dy.data_ptr(), dx.data_ptr(), x.numel(), _dt(x), _stream()))
        return dx


def gelu(x):
    return _GELU.apply(x)


class _CausalAttention(Function):
    """Causal self-attention (head dim 64) straight on the fused projection: ``qkv`` [B*T, 3*H*64] bf16 -> [B*T, H*64] bf16.
    Flash attention on tcgen05 / TMEM (csrc/ops/attention_tc.cu): q / k / v are read in place through one TMA tensor map, the
    output comes back with the heads already merged, and backward returns d(qkv) in the projection's own layout - so the
    split-heads / merge-heads copies of the module-level formulation disappear together with the library SDPA call."""

    @staticmethod
    def forward(ctx, qkv, B, T, H):
        qkv = qkv if qkv.is_contiguous() else qkv.contiguous()
        out = torch.empty(B * T, H * 64, dtype=torch.bfloat16, device=qkv.device)
        lse = torch.empty(B * H, T, dtype=torch.float32, device=qkv.device)
        _chk(lib().dcp_attn_fwd(qkv.data_ptr(), out.data_ptr(), lse.data_ptr(), B, T, H, _stream()))
        ctx.save_for_backward(qkv, out, lse)
        ctx.dims = (B, T, H)
        return out

    @staticmethod
    def backward(ctx, dout):
        qkv, out, lse = ctx.saved_tensors
        B, T, H = ctx.dims
        dout = dout if dout.is_contiguous() else dout.contiguous()
        dqkv = torch.empty_like(qkv)
        delta = torch.empty(B * H * T, dtype=torch.float32, device=qkv.device)
        acc = torch.empty(B * T * H * 64, dtype=torch.float32, device=qkv.device)
        _chk(lib().dcp_attn_bwd(qkv.data_ptr(), out.data_ptr(), dout.data_ptr(), lse.data_ptr(), dqkv.data_ptr(), delta.data_ptr(), acc.data_ptr(),
                                B, T, H, _stream()))
        return dqkv, None, None, None


def causal_attention_supported(qkv: torch.Tensor, T: int, H: int) -> bool:
    return qkv.is_cuda and qkv.dtype == torch.bfloat16 and T % 128 == 0 and qkv.shape[-1] == 3 * H * 64


def causal_attention(qkv: torch.Tensor, B: int, T: int, H: int) -> torch.Tensor:
    return _CausalAttention.apply(qkv, B, T, H)


# -------------------------------------------------------------------------------------- optimizer --
def _push_hyper(opt, group: dict, ent: dict) -> None:
    hyper = opt._hyper(group)
    if ent["hyper"] != hyper:
        vals = (hyper + [0.0] * 6)[:6]
        ent["hp"][:6].copy_(torch.tensor(vals, dtype=torch.float32), non_blocking=False)
        ent["hp"][7] = 1.0
        ent["hyper"] = list(hyper)


def sync_optimizer_hyper(opt) -> None:
    """Bring the device-resident hyper-parameter blocks in line with ``opt.param_groups`` (e.g. after ``StepLR.step()``).
    The captured training graphs call this before every replay: the kernels inside read lr & co. from device memory, so
    a changed learning rate takes effect without re-capturing.  A no-op (a Python list compare) when nothing changed."""
    cache = opt.__dict__.get("_dcp_cache")
    if not cache:
        return
    for gi, group in enumerate(opt.param_groups):
        ent = cache.get(gi)
        if ent is not None and ent["hyper"] is not None and ent["hyper"] != opt._hyper(group):
            _push_hyper(opt, group, ent)


def fused_optimizer_step(opt, gi: int, group: dict, items: List[Tuple[torch.Tensor, torch.Tensor, dict]]) -> None:
    """One multi-tensor launch for ``items`` = [(param, grad, state)].  The pointer table and the
    hyper-parameter block live on the device and are rebuilt only when pointers / values change."""
    cache = opt.__dict__.setdefault("_dcp_cache", {})
    ent = cache.get(gi)
    dev = items[0][0].device
    gdt = items[0][1].dtype
    # the table bakes in parameter, gradient AND state pointers: load_state_dict() replaces the state tensors, so they are part of the key
    def _key():
        return tuple((p.data_ptr(), p.grad.data_ptr()) + tuple(t.data_ptr() for t in opt._state_tensors(st)) for p, _, st in items)
    key = _key()
    if ent is None or ent["key"] != key:
        rows = []
        for p, g, st in items:
            if p.dtype != torch.float32:
                raise TypeError("fused optimizers keep parameters in float32")
            if g.dtype != gdt:
                raise TypeError("all gradients of a param group must share a dtype")
            if g.stride() != p.stride():
                g2 = torch.empty_like(p, dtype=g.dtype)
                g2.copy_(g)
                p.grad = g = g2
            sts = opt._state_tensors(st)
            for s_ in sts:
                if s_.stride() != p.stride():
                    raise RuntimeError("optimizer state layout differs from its parameter")
            s1 = sts[0].data_ptr() if len(sts) > 0 else 0
            s2 = sts[1].data_ptr() if len(sts) > 1 else 0
            rows.append(struct.pack("QQQQq", p.data_ptr(), g.data_ptr(), s1, s2, p.numel()))
        key = _key()
        # pinned staging + async copy: legal inside CUDA-graph capture (becomes a memcpy node)
        host = torch.frombuffer(bytearray(b"".join(rows)), dtype=torch.uint8).pin_memory()
        table = torch.empty(host.numel(), dtype=torch.uint8, device=dev)
        table.copy_(host, non_blocking=True)
        first_step = items[0][2].get("step", 1)
        hp = ent["hp"] if ent is not None else torch.zeros(8, dtype=torch.float32, device=dev)
        # (re)seed the device step counter from the optimizer state: first build, or a rebuild because load_state_dict() replaced
        # the state tensors (a rebuild for new gradient pointers - e.g. under graph capture - keeps the running counter)
        if ent is None or [k[2:] for k in ent["key"]] != [k[2:] for k in key]:
            hp[6] = float(int(first_step) - 1)
        ent = {"key": key, "table": table, "host": host, "hp": hp, "hyper": ent["hyper"] if ent else None, "n": len(items),
               "max": max(p.numel() for p, _, _ in items)}
        cache[gi] = ent
    _push_hyper(opt, group, ent)
    _chk(lib().dcp_optim_advance(ent["hp"].data_ptr(), _stream()))
    _chk(lib().dcp_optim_step(ent["table"].data_ptr(), ent["n"], ent["max"], ent["hp"].data_ptr(), opt._kind, _DT[gdt], _stream()))
