"""Sharded optimizer whose step is fused with the gradient collective: ONE kernel per bucket does
reduce_scatter -> Adadelta / SGD-momentum / Adam(W) update -> all_gather over NVLink peer memory
(csrc/coll/fused_opt.cu).  Optimizer state is kept only for the 1/W shard a rank owns (ZeRO-1), parameters are
re-homed into flat symmetric buffers that mirror the DDP buckets, and the kernel is launched from the DDP
bucket-ready hook on the side stream - so the update of bucket k overlaps the backward of the layers behind it.

This is the north-star's "fused reduce_scatter -> optimizer -> all_gather step" / "large-bucket gradient
all_reduce fused with optimizer step" (BASELINE.json); the reference itself runs a replicated
``optimizer.step()`` after DDP's all-reduce (ref:main.py:62-63).  ``state_dict()`` of the model still yields
full per-parameter tensors (the views are dense), so checkpoints keep the reference's format.
"""
from __future__ import annotations

import ctypes as C
from typing import List

import torch

from ..ops import cuda as K
from .ddp import DistributedDataParallel
from .p2p_backend import CH_OPT

_KINDS = {"adadelta": 0, "sgd": 1, "adam": 2, "adamw": 2}


class FusedShardedOptimizer:
    def __init__(self, ddp: DistributedDataParallel, kind: str = "adam", lr: float = 1e-3, rho: float = 0.9, eps: float = None,
                 momentum: float = 0.0, dampening: float = 0.0, nesterov: bool = False, betas=(0.9, 0.999), weight_decay: float = 0.0,
                 use_nvls: bool = True, max_ctas: int = 0, shadow_dtype=None):
        if ddp._cuda_backend is None:
            raise RuntimeError("FusedShardedOptimizer needs the NVLink P2P backend (CUDA parameters, world_size >= 1)")
        self.ddp, self.be = ddp, ddp._cuda_backend
        self.kind_name, self.kind = kind.lower(), _KINDS[kind.lower()]
        self.lr, self.use_nvls, self.max_ctas = lr, use_nvls, max_ctas
        if self.kind == 0:
            self._hyper = lambda: [self.lr, rho, 1e-6 if eps is None else eps, weight_decay, 0.0, 0.0]
        elif self.kind == 1:
            self._hyper = lambda: [self.lr, momentum, dampening, weight_decay, 1.0 if nesterov else 0.0, 0.0]
        else:
            self._hyper = lambda: [self.lr, betas[0], betas[1], 1e-8 if eps is None else eps, weight_decay, 1.0 if self.kind_name == "adamw" else 0.0]
        lib = K.lib()
        cl = self.be.lib
        cl.dcp_fused_shard_elems.restype = C.c_size_t
        cl.dcp_fused_shard_elems.argtypes = [C.c_size_t, C.c_int, C.c_int]
        cl.dcp_fused_rs_opt_ag.restype = C.c_int
        cl.dcp_fused_rs_opt_ag.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_int, C.c_int,
                                           C.c_int, C.c_int, C.c_int, C.c_void_p]
        self._lib = lib
        if not ddp._rebuilt:
            ddp.finalize_buckets()
        dev = ddp.device
        self.hp = torch.zeros(8, dtype=torch.float32, device=dev)
        self._last_hyper = None
        self.param_bufs: List[torch.Tensor] = []
        # shadow_dtype (e.g. torch.bfloat16): keep a low-precision copy of every flat parameter buffer, refreshed by ONE cast launch
        # per bucket right behind the update; the Linear layers multiply with it instead of casting their weight every forward
        self.shadow_dtype = shadow_dtype
        self.shadow_bufs: List[torch.Tensor] = []
        self.states = []
        nstate = 1 if self.kind == 1 else 2
        for b in ddp._buckets:
            n = b.flat.numel()
            npad = -(-n // 8) * 8
            pbuf = self.be.empty_symmetric(npad, torch.float32)
            pbuf.zero_()
            # re-home the parameters: same offsets as their gradient views inside the bucket
            base = b.flat.data_ptr()
            for p, v in zip(b.params, b.views):
                if p.dtype != torch.float32:
                    raise TypeError("FusedShardedOptimizer keeps fp32 master parameters")
                off = (v.data_ptr() - base) // b.flat.element_size()
                dst = pbuf[off: off + p.numel()]
                dst = dst.view(p.shape) if p.is_contiguous() else dst.as_strided(p.shape, p.stride())
                dst.copy_(p.detach())
                p.data = dst
            if shadow_dtype is not None:
                sbuf = torch.empty(npad, dtype=shadow_dtype, device=dev)
                for p, v in zip(b.params, b.views):
                    if p.is_contiguous():
                        off = (v.data_ptr() - base) // b.flat.element_size()
                        K.register_shadow(p.data, sbuf[off: off + p.numel()].view(p.shape))
                self.shadow_bufs.append(sbuf)
            gd = K._DT[b.flat.dtype]
            shard = int(cl.dcp_fused_shard_elems(n, self.be.world, gd))
            self.param_bufs.append(pbuf)
            self.states.append([torch.zeros(shard, dtype=torch.float32, device=dev) for _ in range(nstate)])
        ddp._bucket_override = self._on_bucket_ready
        self._sync_hyper()
        self.refresh_shadows()
        if self.shadow_bufs:
            # parameters written through load_state_dict() must reach the shadows too (in-place writes the module system does not
            # see - p.data.copy_() by hand - need an explicit refresh_shadows())
            self._load_hook = ddp.module.register_load_state_dict_post_hook(lambda module, incompatible: self.refresh_shadows())

    def refresh_shadows(self) -> None:
        """Re-cast every flat parameter buffer into its low-precision shadow (after loading a checkpoint into the parameters)."""
        for pbuf, sbuf in zip(self.param_bufs, self.shadow_bufs):
            K._chk(self._lib.dcp_cast_from_f32(pbuf.data_ptr(), sbuf.data_ptr(), pbuf.numel(), K._DT[sbuf.dtype], K._stream()))

    def close(self) -> None:
        if getattr(self, "_load_hook", None) is not None:
            self._load_hook.remove()
            self._load_hook = None
        if self.shadow_bufs:
            K.drop_shadows([p.data for b in self.ddp._buckets for p in b.params])
            self.shadow_bufs = []

    def _sync_hyper(self) -> None:
        h = self._hyper()
        if h != self._last_hyper:
            self.hp[:6].copy_(torch.tensor(h, dtype=torch.float32))
            self.hp[7] = 1.0
            self._last_hyper = list(h)

    def _on_bucket_ready(self, bucket) -> None:
        """Called by DDP (instead of the all-reduce) as soon as every gradient of ``bucket`` has landed."""
        be = self.be
        cur = torch.cuda.current_stream(be._torch_dev)
        be.comm_stream.wait_stream(cur)
        i = bucket.index
        with torch.cuda.stream(be.comm_stream):
            if i == 0:
                K._chk(self._lib.dcp_optim_advance(self.hp.data_ptr(), be.comm_stream.cuda_stream))
            st = self.states[i]
            rc = be.lib.dcp_fused_rs_opt_ag(be.handle, bucket.flat.data_ptr(), self.param_bufs[i].data_ptr(), st[0].data_ptr(),
                                           st[1].data_ptr() if len(st) > 1 else None, bucket.flat.numel(), self.hp.data_ptr(), self.kind,
                                           K._DT[bucket.flat.dtype], int(self.use_nvls), CH_OPT, self.max_ctas, be.comm_stream.cuda_stream)
            if rc == 0 and self.shadow_bufs:
                sb = self.shadow_bufs[i]
                rc = self._lib.dcp_cast_from_f32(self.param_bufs[i].data_ptr(), sb.data_ptr(), sb.numel(), K._DT[sb.dtype], be.comm_stream.cuda_stream)
        be._check(rc, "fused reduce_scatter/optimizer/all_gather")
        ev = torch.cuda.Event()
        ev.record(be.comm_stream)
        bucket.work_event = ev

    # -- torch.optim-like surface ---------------------------------------------------------------------
    @property
    def param_groups(self):
        return [{"lr": self.lr, "params": [p for b in self.ddp._buckets for p in b.params]}]

    def set_lr(self, lr: float) -> None:
        self.lr = lr
        self._sync_hyper()

    def zero_grad(self, set_to_none: bool = True) -> None:
        for p in self.ddp._params:
            if set_to_none:
                p.grad = None
            elif p.grad is not None:
                p.grad.zero_()

    def step(self) -> None:
        """The update already ran bucket by bucket during backward; this only orders it before later work."""
        for b in self.ddp._buckets:
            if b.work_event is not None:
                torch.cuda.current_stream().wait_event(b.work_event)
                b.work_event = None
