"""Optimizers and LR schedule.

``Adadelta`` + ``StepLR`` are what the reference trains with (ref:main.py:124-125,63,131); SGD-momentum
and Adam are the north-star additions.  On CUDA every optimizer is ONE multi-tensor kernel launch per
step (csrc/ops/optim.cu) that walks a device-resident table of (param, grad, state...) pointers; the
learning rate lives in device memory so a captured CUDA graph keeps working when ``StepLR`` changes it.
The CPU path is the same math in torch ops.  ``state_dict()`` layouts follow torch.optim.
"""
from __future__ import annotations

from typing import List

import torch
from torch.optim import Optimizer


class _FusedBase(Optimizer):
    _kind = -1  # kernel selector
    _nstate = 0

    def _init_state(self, p: torch.Tensor) -> dict:
        raise NotImplementedError

    def _state_tensors(self, st: dict) -> List[torch.Tensor]:
        raise NotImplementedError

    def _hyper(self, group: dict) -> List[float]:
        raise NotImplementedError

    def _cpu_update(self, p, g, st, group) -> None:
        raise NotImplementedError

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        for gi, group in enumerate(self.param_groups):
            cuda_items = []
            for p in group["params"]:
                if p.grad is None:
                    continue
                st = self.state[p]
                if not st:
                    st.update(self._init_state(p))
                if "step" in st:
                    if torch.is_tensor(st["step"]):
                        if not p.is_cuda:
                            st["step"] += 1
                    else:
                        st["step"] += 1
                if p.is_cuda:
                    cuda_items.append((p, p.grad, st))
                else:
                    self._cpu_update(p, p.grad, st, group)
            if cuda_items:
                from . import cuda
                cuda.fused_optimizer_step(self, gi, group, cuda_items)
        return loss


class Adadelta(_FusedBase):
    """``Adadelta(params, lr=1.0, rho=0.9, eps=1e-6, weight_decay=0)`` (the reference passes lr=0.001)."""
    _kind, _nstate = 0, 2

    def __init__(self, params, lr: float = 1.0, rho: float = 0.9, eps: float = 1e-6, weight_decay: float = 0.0):
        super().__init__(params, dict(lr=lr, rho=rho, eps=eps, weight_decay=weight_decay))

    def _init_state(self, p):
        return {"step": 0, "square_avg": torch.zeros_like(p), "acc_delta": torch.zeros_like(p)}

    def _state_tensors(self, st):
        return [st["square_avg"], st["acc_delta"]]

    def _hyper(self, g):
        return [g["lr"], g["rho"], g["eps"], g["weight_decay"]]

    def _cpu_update(self, p, g, st, group):
        lr, rho, eps, wd = self._hyper(group)
        if wd:
            g = g.add(p, alpha=wd)
        sq, acc = st["square_avg"], st["acc_delta"]
        sq.mul_(rho).addcmul_(g, g, value=1 - rho)
        std = sq.add(eps).sqrt_()
        delta = acc.add(eps).sqrt_().div_(std).mul_(g)
        acc.mul_(rho).addcmul_(delta, delta, value=1 - rho)
        p.add_(delta, alpha=-lr)


class SGD(_FusedBase):
    _kind, _nstate = 1, 1

    def __init__(self, params, lr: float = 1e-3, momentum: float = 0.0, dampening: float = 0.0, weight_decay: float = 0.0,
                 nesterov: bool = False):
        super().__init__(params, dict(lr=lr, momentum=momentum, dampening=dampening, weight_decay=weight_decay, nesterov=nesterov))

    def _init_state(self, p):
        return {"step": 0, "momentum_buffer": torch.zeros_like(p)}

    def _state_tensors(self, st):
        return [st["momentum_buffer"]]

    def _hyper(self, g):
        return [g["lr"], g["momentum"], g["dampening"], g["weight_decay"], 1.0 if g["nesterov"] else 0.0]

    def _cpu_update(self, p, g, st, group):
        lr, mom, damp, wd, nest = self._hyper(group)
        if wd:
            g = g.add(p, alpha=wd)
        if mom:
            buf = st["momentum_buffer"]
            if st["step"] == 1:
                buf.copy_(g)
            else:
                buf.mul_(mom).add_(g, alpha=1 - damp)
            g = g.add(buf, alpha=mom) if nest else buf
        p.add_(g, alpha=-lr)


class Adam(_FusedBase):
    """Adam / AdamW (``decoupled=True``)."""
    _kind, _nstate = 2, 2

    def __init__(self, params, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8, weight_decay: float = 0.0,
                 decoupled: bool = False):
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, decoupled=decoupled))

    def _init_state(self, p):
        return {"step": 0, "exp_avg": torch.zeros_like(p, dtype=torch.float32), "exp_avg_sq": torch.zeros_like(p, dtype=torch.float32)}

    def _state_tensors(self, st):
        return [st["exp_avg"], st["exp_avg_sq"]]

    def _hyper(self, g):
        return [g["lr"], g["betas"][0], g["betas"][1], g["eps"], g["weight_decay"], 1.0 if g["decoupled"] else 0.0]

    def _cpu_update(self, p, g, st, group):
        lr, b1, b2, eps, wd, dec = self._hyper(group)
        t = st["step"]
        if wd:
            if dec:
                p.mul_(1 - lr * wd)
            else:
                g = g.add(p, alpha=wd)
        m, v = st["exp_avg"], st["exp_avg_sq"]
        m.mul_(b1).add_(g, alpha=1 - b1)
        v.mul_(b2).addcmul_(g, g, value=1 - b2)
        bc1, bc2 = 1 - b1 ** t, 1 - b2 ** t
        p.addcdiv_(m, (v / bc2).sqrt_().add_(eps), value=-lr / bc1)


AdamW = lambda params, **kw: Adam(params, decoupled=True, **kw)  # noqa: E731


class StepLR:
    """``lr = base_lr * gamma ** (epoch // step_size)`` (ref:main.py:125,131: step_size=1, gamma=0.7)."""

    def __init__(self, optimizer: Optimizer, step_size: int, gamma: float = 0.1, last_epoch: int = -1):
        self.optimizer, self.step_size, self.gamma = optimizer, step_size, gamma
        self.base_lrs = [g.setdefault("initial_lr", g["lr"]) for g in optimizer.param_groups]
        self.last_epoch = last_epoch
        self.step()

    def get_last_lr(self):
        return [g["lr"] for g in self.optimizer.param_groups]

    def step(self) -> None:
        self.last_epoch += 1
        for g, base in zip(self.optimizer.param_groups, self.base_lrs):
            g["lr"] = base * self.gamma ** (self.last_epoch // self.step_size)

    def state_dict(self):
        return {k: v for k, v in self.__dict__.items() if k != "optimizer"}

    def load_state_dict(self, sd):
        self.__dict__.update(sd)
