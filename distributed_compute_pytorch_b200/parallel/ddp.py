"""DistributedDataParallel: replicated module, bucketed gradient all-reduce overlapped with backward.

What the reference gets from ``DistributedDataParallel(model, device_ids=[rank])`` (ref:main.py:122):

* the wrapped network lives at ``.module`` so ``state_dict()`` keys carry the ``module.`` prefix that
  ends up in ``mnist.pt`` (ref:main.py:133);
* construction checks that every rank holds the same parameter shapes and broadcasts rank 0's
  parameters and buffers;
* every training forward re-broadcasts the buffers (BatchNorm running stats) from rank 0;
* ``loss.backward()`` (ref:main.py:62) averages gradients across ranks bucket by bucket while
  autograd is still running; the first iteration uses one bucket, later ones are rebuilt in observed
  gradient-ready order with a 1 MiB first bucket and a 25 MiB cap.

B200 design: buckets are carved out of the NVLink symmetric heap, so the hand-written P2P all-reduce
(csrc/coll) reads peers' buckets directly - no staging copy - and fuses the 1/world_size average and
the bucket-dtype cast into the reduction.  Bucket collectives run on a dedicated side stream with a
capped CTA count so backward GEMMs keep their SMs.  Ready tracking / bucket planning is native
(csrc/ddp/reducer.cpp).
"""
from __future__ import annotations

import ctypes as C
import hashlib
from contextlib import contextmanager
from typing import List, Optional

import torch
from torch import nn

from .. import distributed as dist
from .._native import host_lib
from ..types import ReduceOp

FIRST_BUCKET_BYTES = 1024 * 1024
DEFAULT_BUCKET_CAP_MB = 25
BROADCAST_BUCKET_BYTES = 250 * 1024 * 1024


def plan_buckets(nbytes: List[int], group_keys: List[int], limits: List[int]) -> List[int]:
    """bucket index per tensor, tensors given in (expected) gradient-ready order."""
    lib = host_lib()
    n = len(nbytes)
    if n == 0:
        return []
    a = (C.c_int64 * n)(*nbytes)
    k = (C.c_int32 * n)(*group_keys)
    lim = (C.c_int64 * len(limits))(*limits)
    out = (C.c_int32 * n)()
    lib.dcp_bucket_assign(a, k, n, lim, len(limits), out)
    return list(out)


def broadcast_coalesced(tensors: List[torch.Tensor], src: int = 0, buffer_size: int = BROADCAST_BUCKET_BYTES) -> None:
    """Broadcast many tensors from ``src`` with one collective per (dtype, device) chunk."""
    groups = {}
    for t in tensors:
        groups.setdefault((t.dtype, t.device), []).append(t)
    for (_dtype, _dev), ts in groups.items():
        chunk, size = [], 0
        def flush():
            if not chunk:
                return
            if len(chunk) == 1 and chunk[0].is_contiguous():
                dist.broadcast(chunk[0].detach(), src)
            else:
                flat = torch.cat([t.detach().reshape(-1) for t in chunk])
                dist.broadcast(flat, src)
                off = 0
                for t in chunk:
                    t.detach().copy_(flat[off: off + t.numel()].view(t.shape))
                    off += t.numel()
        for t in ts:
            nb = t.numel() * t.element_size()
            if chunk and size + nb > buffer_size:
                flush()
                chunk, size = [], 0
            chunk.append(t)
            size += nb
        flush()


class _Bucket:
    __slots__ = ("flat", "params", "views", "index", "work_event")

    def __init__(self, flat, params, views, index):
        self.flat, self.params, self.views, self.index, self.work_event = flat, params, views, index, None


class GradBucket:
    """Argument handed to communication hooks (``register_comm_hook``)."""

    def __init__(self, bucket: _Bucket, is_last: bool):
        self._b, self._last = bucket, is_last

    def buffer(self) -> torch.Tensor: return self._b.flat
    def index(self) -> int: return self._b.index
    def is_last(self) -> bool: return self._last
    def parameters(self): return list(self._b.params)
    def gradients(self): return list(self._b.views)


class DistributedDataParallel(nn.Module):
    def __init__(self, module: nn.Module, device_ids=None, output_device=None, dim: int = 0, broadcast_buffers: bool = True,
                 init_sync: bool = True, process_group=None, bucket_cap_mb: Optional[float] = None,
                 find_unused_parameters: bool = False, check_reduction: bool = False, gradient_as_bucket_view: bool = False,
                 static_graph: bool = False, grad_dtype: Optional[torch.dtype] = None, comm_ctas: Optional[int] = None):
        super().__init__()
        if not dist.is_initialized():
            raise RuntimeError("Default process group has not been initialized, please make sure to call init_process_group.")
        self.module = module
        self.device_ids, self.output_device, self.dim = device_ids, output_device, dim
        self.broadcast_buffers, self.find_unused_parameters = broadcast_buffers, find_unused_parameters
        self.gradient_as_bucket_view, self.static_graph = gradient_as_bucket_view, static_graph
        self.bucket_bytes_cap = int((DEFAULT_BUCKET_CAP_MB if bucket_cap_mb is None else bucket_cap_mb) * 1024 * 1024)
        # the small first bucket only applies with the default cap (same rule as the library wrapper)
        self.first_bucket_bytes = FIRST_BUCKET_BYTES if bucket_cap_mb is None else self.bucket_bytes_cap
        self.grad_dtype = grad_dtype
        self.comm_ctas = comm_ctas
        self.world_size, self.rank = dist.get_world_size(), dist.get_rank()
        self.require_backward_grad_sync = True
        self.require_forward_param_sync = True
        self._comm_hook = None
        self._params = [p for p in module.parameters() if p.requires_grad]
        if not self._params:
            raise RuntimeError("DistributedDataParallel is not needed when a module doesn't have any parameter that requires a gradient.")
        devs = {p.device for p in self._params}
        if len(devs) != 1:
            raise ValueError("DistributedDataParallel requires all parameters on one device")
        self.device = next(iter(devs))
        self._pg = dist._world(process_group)
        self._cuda_backend = self._pg.cuda if self.device.type == "cuda" else None
        self._verify_params_across_processes()
        if init_sync:
            self._sync_module_states()
        # iteration 1: a single unbounded bucket per dtype, in reverse registration order
        order = list(range(len(self._params)))[::-1]
        self._buckets: List[_Bucket] = []
        self._reducer = None
        self._build_buckets(order, limits=[2 ** 62])
        self._rebuilt = False
        self._iter_callback_queued = False
        self._hooks = [p.register_post_accumulate_grad_hook(self._make_hook(i)) for i, p in enumerate(self._params)]
        self._lib = host_lib()

    # ------------------------------------------------------------------ construction helpers ----
    def _verify_params_across_processes(self) -> None:
        desc = ";".join(f"{tuple(p.shape)}:{p.dtype}" for p in self._params)
        digest = hashlib.sha1(desc.encode()).hexdigest()
        seq = getattr(self._pg, "_ddp_count", 0)
        self._pg._ddp_count = seq + 1
        key = f"ddp_verify/{seq}"
        store = self._pg.pstore
        if self.rank == 0:
            store.set(key + "/0", digest)
        ref = store.get(key + "/0").decode()
        if ref != digest:
            raise RuntimeError(
                f"DDP expects same model across all ranks, but rank {self.rank} has {len(self._params)} params with a "
                "different shape/dtype signature than rank 0")
        store.barrier(key + "/b", self.rank, self.world_size)

    def _sync_module_states(self) -> None:
        tensors = [p.detach() for p in self.module.parameters()] + [b.detach() for b in self.module.buffers()]
        broadcast_coalesced(tensors, 0)

    def _alloc_flat(self, numel: int, dtype: torch.dtype) -> torch.Tensor:
        if self._cuda_backend is not None:
            return self._cuda_backend.empty_symmetric(numel, dtype)
        return torch.zeros(numel, dtype=dtype, device=self.device)

    def _build_buckets(self, order: List[int], limits: List[int]) -> None:
        """(Re)plan buckets for parameters visited in ``order`` and allocate their flat storage."""
        for b in self._buckets:
            if self._cuda_backend is not None:
                self._cuda_backend.free_symmetric(b.flat)
        dtypes = {}
        nbytes, keys = [], []
        for i in order:
            p = self._params[i]
            dt = self.grad_dtype or p.dtype
            nbytes.append(p.numel() * torch.empty((), dtype=dt).element_size())
            keys.append(dtypes.setdefault(dt, len(dtypes)))
        assign = plan_buckets(nbytes, keys, limits)
        nb = max(assign) + 1 if assign else 0
        members: List[List[int]] = [[] for _ in range(nb)]
        for pos, i in enumerate(order):
            members[assign[pos]].append(i)
        self._buckets = []
        self._bucket_of = [0] * len(self._params)
        self._view_of: List[Optional[torch.Tensor]] = [None] * len(self._params)
        for bi, idxs in enumerate(members):
            dt = self.grad_dtype or self._params[idxs[0]].dtype
            # 16-byte aligned views so the vectorised kernels can use 128-bit accesses on each slice
            align = max(1, 16 // torch.empty((), dtype=dt).element_size())
            offs, total = [], 0
            for i in idxs:
                offs.append(total)
                total += -(-self._params[i].numel() // align) * align
            flat = self._alloc_flat(total, dt)
            flat.zero_()
            views = []
            for i, off in zip(idxs, offs):
                p = self._params[i]
                # same physical layout as the parameter (e.g. channels-last conv weights): kernels write the
                # gradient in that order and the fused optimizer walks param / grad / state with one index
                if p.is_contiguous():
                    v = flat[off: off + p.numel()].view(p.shape)
                else:
                    v = flat[off: off + p.numel()].as_strided(p.shape, p.stride())
                views.append(v)
                self._bucket_of[i] = bi
                self._view_of[i] = v
            self._buckets.append(_Bucket(flat, [self._params[i] for i in idxs], views, bi))
        lib = host_lib()
        if self._reducer is not None:
            lib.dcp_reducer_destroy(self._reducer)
        arr = (C.c_int32 * len(self._bucket_of))(*self._bucket_of)
        self._reducer = lib.dcp_reducer_create(arr, len(self._bucket_of), nb)

    def bucket_sizes_bytes(self) -> List[int]:
        return [sum(p.numel() * b.flat.element_size() for p in b.params) for b in self._buckets]

    # ------------------------------------------------------------------------- backward path ----
    def _make_hook(self, idx: int):
        def hook(param: torch.Tensor) -> None:
            if not self.require_backward_grad_sync or (self.world_size == 1 and self._comm_hook is None
                                                        and getattr(self, "_bucket_override", None) is None):
                return
            if not self._iter_callback_queued:
                self._iter_callback_queued = True
                torch.autograd.Variable._execution_engine.queue_callback(self._finalize_backward)
            view = self._view_of[idx]
            g = param.grad
            if g is not None and g.data_ptr() != view.data_ptr():
                view.copy_(g)  # cast to the bucket dtype happens here when they differ
                if view.dtype == param.dtype:
                    param.grad = view  # alias: the reduced value is visible without a copy-back
            first = C.c_int(0)
            k = self._lib.dcp_reducer_mark_ready(self._reducer, idx, C.byref(first))
            if k < 0:
                raise RuntimeError("a parameter was marked ready twice in one backward; use no_sync() for gradient accumulation")
            for b in range(first.value, first.value + k):
                self._launch_bucket(self._buckets[b])
        return hook

    def finalize_buckets(self, order: Optional[List[int]] = None) -> None:
        """Fix the bucket layout now (capped buckets in ``order``, default reverse registration order) instead of
        waiting for the first backward - used by engines that need stable bucket addresses up front."""
        order = list(range(len(self._params)))[::-1] if order is None else order
        for i, p in enumerate(self._params):
            if p.grad is not None and self._view_of[i] is not None and p.grad.data_ptr() == self._view_of[i].data_ptr():
                p.grad = None
        self._build_buckets(order, limits=[self.first_bucket_bytes, self.bucket_bytes_cap])
        self._rebuilt = True
        self._pending_order = None

    def _launch_bucket(self, bucket: _Bucket) -> None:
        if getattr(self, "_bucket_override", None) is not None:
            self._bucket_override(bucket)
            return
        if self._comm_hook is not None:
            state, fn = self._comm_hook
            out = fn(state, GradBucket(bucket, bucket.index == len(self._buckets) - 1))
            if hasattr(out, "wait"):
                out = out.wait()
            if isinstance(out, (list, tuple)):
                out = out[0]
            if out is not None and out.data_ptr() != bucket.flat.data_ptr():
                bucket.flat.copy_(out)
            return
        if self._cuda_backend is not None:
            timing = getattr(self, "_comm_timing", None)
            if timing is not None:      # opt-in (enable_comm_timing): CUDA events on the side stream around this bucket's collective
                be = self._cuda_backend
                be.comm_stream.wait_stream(torch.cuda.current_stream(be._torch_dev))
                e0 = torch.cuda.Event(enable_timing=True); e0.record(be.comm_stream)
            bucket.work_event = self._cuda_backend.all_reduce_async(bucket.flat, ReduceOp.AVG, max_ctas=self.comm_ctas)
            if timing is not None:
                e1 = torch.cuda.Event(enable_timing=True); e1.record(be.comm_stream)
                timing.append((bucket.index, bucket.flat.numel() * bucket.flat.element_size(), e0, e1))
        else:
            dist.all_reduce(bucket.flat, ReduceOp.AVG)

    def _finalize_backward(self) -> None:
        self._iter_callback_queued = False
        lib = self._lib
        if lib.dcp_reducer_unlaunched(self._reducer) > 0:
            if not self.find_unused_parameters:
                lib.dcp_reducer_reset(self._reducer)
                raise RuntimeError(
                    "Expected to have finished reduction in the prior iteration: some parameters did not receive a "
                    "gradient. Pass find_unused_parameters=True to DistributedDataParallel.")
            n = len(self._params)
            order = (C.c_int32 * n)()
            got = lib.dcp_reducer_ready_order(self._reducer, order, n)
            seen = set(order[:got])
            for i in range(n):
                if i not in seen:
                    self._view_of[i].zero_()
                    first = C.c_int(0)
                    k = lib.dcp_reducer_mark_ready(self._reducer, i, C.byref(first))
                    for b in range(first.value, first.value + k):
                        self._launch_bucket(self._buckets[b])
        for b in self._buckets:
            if b.work_event is not None:
                torch.cuda.current_stream().wait_event(b.work_event)
                b.work_event = None
        # bucket dtype differs from param dtype (e.g. bf16 buckets for fp32 params): copy the average back
        for i, p in enumerate(self._params):
            v = self._view_of[i]
            if v.dtype != p.dtype and p.grad is not None:
                p.grad.copy_(v)
        if not self._rebuilt and not self.static_graph:
            n = len(self._params)
            order = (C.c_int32 * n)()
            got = lib.dcp_reducer_ready_order(self._reducer, order, n)
            self._pending_order = list(order[:got]) + [i for i in range(n) if i not in set(order[:got])]
        lib.dcp_reducer_reset(self._reducer)

    def _maybe_rebuild_buckets(self) -> None:
        order = getattr(self, "_pending_order", None)
        if order is None or self._rebuilt:
            return
        self._rebuilt = True
        self._pending_order = None
        # gradients that alias the old buckets must survive the re-allocation
        for i, p in enumerate(self._params):
            if p.grad is not None and p.grad.data_ptr() == self._view_of[i].data_ptr():
                p.grad = p.grad.clone()
        self._build_buckets(order, limits=[self.first_bucket_bytes, self.bucket_bytes_cap])

    # -------------------------------------------------------------------------- forward path ----
    def _sync_buffers(self) -> None:
        bufs = [b for b in self.module.buffers()]
        if bufs:
            with torch.no_grad():
                broadcast_coalesced(bufs, 0)

    def enable_comm_timing(self, on: bool = True) -> None:
        """Record CUDA events around every gradient bucket's all-reduce; read them with ``comm_timings()``.
        Not for use inside a captured CUDA graph."""
        self._comm_timing = [] if on else None

    def comm_timings(self, clear: bool = True):
        """[(bucket index, bytes, milliseconds on the side stream)] since the last call (synchronises)."""
        rec = getattr(self, "_comm_timing", None) or []
        if rec:
            torch.cuda.synchronize()
        out = [(i, nbytes, e0.elapsed_time(e1)) for i, nbytes, e0, e1 in rec]
        if clear and getattr(self, "_comm_timing", None) is not None:
            self._comm_timing = []
        return out

    def forward(self, *inputs, **kwargs):
        nvtx = torch.cuda.is_available() and torch.cuda.is_initialized()
        if nvtx:
            torch.cuda.nvtx.range_push("DistributedDataParallel.forward")
        try:
            return self._forward(*inputs, **kwargs)
        finally:
            if nvtx:
                torch.cuda.nvtx.range_pop()

    def _forward(self, *inputs, **kwargs):
        if self.world_size > 1:
            if torch.is_grad_enabled() and self.require_backward_grad_sync:
                self._maybe_rebuild_buckets()
            if self.broadcast_buffers and self.require_forward_param_sync:
                self._sync_buffers()
            if torch.is_grad_enabled() and self.require_backward_grad_sync:
                self.require_forward_param_sync = True
            else:
                self.require_forward_param_sync = False
        return self.module(*inputs, **kwargs)

    @contextmanager
    def no_sync(self):
        """Skip gradient synchronisation inside the block (gradient accumulation)."""
        old = self.require_backward_grad_sync
        self.require_backward_grad_sync = False
        try:
            yield
        finally:
            self.require_backward_grad_sync = old

    def register_comm_hook(self, state, hook) -> None:
        """``hook(state, GradBucket) -> Tensor | Work-like``; result replaces the bucket contents."""
        self._comm_hook = (state, hook)

    def __del__(self):
        try:
            if self._reducer is not None:
                host_lib().dcp_reducer_destroy(self._reducer)
                self._reducer = None
        except Exception:
            pass
