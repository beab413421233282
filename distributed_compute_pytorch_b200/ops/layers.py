"""nn-style layers.  Parameter names/shapes/initialisation are those of the torch layers the
reference builds at ref:main.py:23-29, so ``state_dict()`` stays interchangeable; the CUDA forward /
backward run the hand-written sm_100a kernels (``ops/cuda.py`` -> csrc/ops), the CPU path is the
plain PyTorch op and only exists for the no-GPU plumbing configuration."""
from __future__ import annotations

import os
from typing import Optional

import torch
import torch.nn.functional as F
from torch import nn


def _k():
    from . import cuda
    return cuda


def use_kernels(x: torch.Tensor) -> bool:
    """True when ``x`` should go to the hand-written kernels: a CUDA tensor, unless DCP_FORCE_TORCH_OPS=1 (the
    stock-PyTorch baseline arm of bench/train_models.py runs the very same module definitions that way)."""
    return x.is_cuda and os.environ.get("DCP_FORCE_TORCH_OPS") != "1"


class Conv2d(nn.Conv2d):
    """3x3 / 1x1 / 7x7 convolution; CUDA path = implicit GEMM on tcgen05 (or direct for tiny C_in),
    optional fused bias + ReLU epilogue.  Weights are kept channels-last (K,R,S,C physical)."""

    def __init__(self, *args, fuse_relu: bool = False, **kwargs):
        super().__init__(*args, **kwargs)
        self.fuse_relu = fuse_relu
        with torch.no_grad():
            self.weight.data = self.weight.data.contiguous(memory_format=torch.channels_last)

    def _apply(self, fn, recurse=True):
        out = super()._apply(fn, recurse)
        with torch.no_grad():
            if self.weight.dim() == 4 and not self.weight.is_contiguous(memory_format=torch.channels_last):
                self.weight.data = self.weight.data.contiguous(memory_format=torch.channels_last)
        return out

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if use_kernels(x):
            return _k().conv2d(x, self.weight, self.bias, self.stride, self.padding, self.fuse_relu)
        y = F.conv2d(x, self.weight, self.bias, self.stride, self.padding, self.dilation, self.groups)
        return F.relu(y) if self.fuse_relu else y


class Linear(nn.Linear):
    def __init__(self, *args, fuse_relu: bool = False, **kwargs):
        super().__init__(*args, **kwargs)
        self.fuse_relu = fuse_relu

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if use_kernels(x):
            return _k().linear(x, self.weight, self.bias, self.fuse_relu)
        y = F.linear(x, self.weight, self.bias)
        return F.relu(y) if self.fuse_relu else y


class _BatchNorm(nn.modules.batchnorm._BatchNorm):
    """BatchNorm with an optional fused ReLU (forward and backward in one kernel each)."""
    _sync = False

    def __init__(self, num_features, eps=1e-5, momentum=0.1, affine=True, track_running_stats=True, fuse_relu: bool = False,
                 **kw):
        super().__init__(num_features, eps, momentum, affine, track_running_stats, **kw)
        self.fuse_relu = fuse_relu

    def _check_input_dim(self, input):
        if input.dim() not in (2, 3, 4):
            raise ValueError(f"expected 2D, 3D or 4D input (got {input.dim()}D input)")

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if use_kernels(x) and x.dim() in (2, 4):
            return _k().batch_norm(x, self.weight, self.bias, self.running_mean, self.running_var, self.num_batches_tracked,
                                   self.training or not self.track_running_stats, self.momentum, self.eps, self.fuse_relu,
                                   self._sync)
        if self._sync and self.training:
            y = _sync_bn_cpu(self, x)
        else:
            y = super().forward(x)
        return F.relu(y) if self.fuse_relu else y


class BatchNorm1d(_BatchNorm):
    pass


class BatchNorm2d(_BatchNorm):
    pass


class SyncBatchNorm(_BatchNorm):
    """BatchNorm whose batch statistics are reduced across ranks (one-shot P2P reduction fused with
    the statistics kernel on CUDA; all_reduce of (sum, sumsq, count) on CPU)."""
    _sync = True

    @classmethod
    def convert_sync_batchnorm(cls, module: nn.Module) -> nn.Module:
        out = module
        if isinstance(module, (nn.modules.batchnorm._BatchNorm,)) and not isinstance(module, SyncBatchNorm):
            out = SyncBatchNorm(module.num_features, module.eps, module.momentum, module.affine, module.track_running_stats,
                                fuse_relu=getattr(module, "fuse_relu", False))
            if module.affine:
                with torch.no_grad():
                    out.weight, out.bias = module.weight, module.bias
            out.running_mean, out.running_var = module.running_mean, module.running_var
            out.num_batches_tracked = module.num_batches_tracked
            out.training = module.training
        for name, child in module.named_children():
            out.add_module(name, cls.convert_sync_batchnorm(child))
        return out


def _sync_bn_cpu(bn: _BatchNorm, x: torch.Tensor) -> torch.Tensor:
    from .. import distributed as dist
    dims = [d for d in range(x.dim()) if d != 1]
    cnt = torch.tensor([float(x.numel() // x.shape[1])])
    s, ss = x.sum(dims), (x * x).sum(dims)
    if dist.is_initialized() and dist.get_world_size() > 1:
        packed = torch.cat([s.detach(), ss.detach(), cnt])
        dist.all_reduce(packed)
        C = x.shape[1]
        # keep the local graph: add the (constant) remote contribution
        s = s + (packed[:C] - s.detach())
        ss = ss + (packed[C:2 * C] - ss.detach())
        cnt = packed[2 * C:]
    mean = s / cnt
    var = ss / cnt - mean * mean
    if bn.track_running_stats:
        with torch.no_grad():
            m = bn.momentum if bn.momentum is not None else 0.1
            bn.running_mean.mul_(1 - m).add_(mean.detach(), alpha=m)
            bn.running_var.mul_(1 - m).add_(var.detach() * (cnt / (cnt - 1).clamp(min=1)), alpha=m)
            bn.num_batches_tracked += 1
    shape = [1, -1] + [1] * (x.dim() - 2)
    y = (x - mean.view(shape)) * torch.rsqrt(var.view(shape) + bn.eps)
    if bn.affine:
        y = y * bn.weight.view(shape) + bn.bias.view(shape)
    return y


class LayerNorm(nn.LayerNorm):
    """LayerNorm over the last dimension (fp32 parameters, f32/bf16 activations)."""

    def forward(self, x):
        if use_kernels(x) and len(self.normalized_shape) == 1 and self.elementwise_affine and self.bias is not None:
            return _k().layer_norm(x, self.weight, self.bias, self.eps)
        return F.layer_norm(x, self.normalized_shape, self.weight, self.bias, self.eps)


class GELU(nn.Module):
    """tanh-approximated GELU (the GPT-2 activation)."""

    def forward(self, x):
        return _k().gelu(x) if use_kernels(x) else F.gelu(x, approximate="tanh")


class ReLU(nn.Module):
    def forward(self, x):
        return _k().relu(x) if use_kernels(x) else F.relu(x)


class MaxPool2d(nn.Module):
    def __init__(self, kernel_size: int, stride: Optional[int] = None, padding: int = 0, out_nchw: bool = False):
        super().__init__()
        self.kernel_size, self.stride, self.padding = kernel_size, stride or kernel_size, padding
        self.out_nchw = out_nchw  # CUDA: write the result NCHW-contiguous so a following flatten(1) is a view

    def forward(self, x):
        if x.is_cuda:
            return _k().max_pool2d(x, self.kernel_size, self.stride, self.padding, self.out_nchw)
        return F.max_pool2d(x, self.kernel_size, self.stride, self.padding)


class Dropout(nn.Module):
    """Element-wise dropout."""

    def __init__(self, p: float = 0.5):
        super().__init__()
        self.p = p

    def forward(self, x):
        if x.is_cuda:
            return _k().dropout(x, self.p, self.training, per_channel=False)
        return F.dropout(x, self.p, self.training)


class Dropout2d(nn.Module):
    """Channel dropout on 4-D input; on a 2-D ``[N, F]`` input it behaves as element-wise dropout,
    which is what the reference's second ``Dropout2d(0.5)`` actually does (ref:main.py:26,42)."""

    def __init__(self, p: float = 0.5):
        super().__init__()
        self.p = p

    def forward(self, x):
        if x.is_cuda:
            return _k().dropout(x, self.p, self.training, per_channel=(x.dim() == 4))
        if x.dim() == 4:
            return F.dropout2d(x, self.p, self.training)
        return F.dropout(x, self.p, self.training)


def log_softmax(x: torch.Tensor, dim: int = 1) -> torch.Tensor:
    if x.is_cuda and x.dim() == 2 and dim in (1, -1):
        return _k().log_softmax(x)
    return F.log_softmax(x, dim=dim)


def nll_loss(logp: torch.Tensor, target: torch.Tensor, reduction: str = "mean") -> torch.Tensor:
    if logp.is_cuda and logp.dim() == 2:
        return _k().nll_loss(logp, target, reduction)
    return F.nll_loss(logp, target, reduction=reduction)


def cross_entropy(logits: torch.Tensor, target: torch.Tensor, reduction: str = "mean") -> torch.Tensor:
    """Fused log_softmax + nll (forward) and (softmax - onehot)/N (backward)."""
    if use_kernels(logits) and logits.dim() == 2:
        return _k().cross_entropy(logits, target, reduction)
    return F.cross_entropy(logits, target, reduction=reduction)


def correct_count(logits: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
    """Number of rows whose argmax equals the label, as a 0-dim int64 tensor (no host sync) - the
    device-side replacement for ``pred.eq(label.view_as(pred)).sum()`` at ref:main.py:83-85."""
    if logits.is_cuda:
        return _k().correct_count(logits, target)
    return logits.argmax(dim=1).eq(target).sum()
