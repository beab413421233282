"""CPU backend: collectives over the native socket mesh (csrc/host/mesh.cpp).

This is the plumbing path - what the reference gets from ``"gloo"`` (ref:main.py:50) - used for CPU
tensors and for the no-GPU configuration (world_size 2 on CPU).  Algorithms are composed here from
full-duplex ``sendrecv``; the reductions themselves run through torch CPU ops so every dtype torch
can add is supported.  Results are bit-identical on all ranks (ring all-gather distributes one copy
of each reduced chunk; the small-message path sums in rank order everywhere).
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional

import torch

from ._native import host_lib
from .types import ReduceOp, Work

_SMALL_BYTES = 32 * 1024


def _apply(op: ReduceOp, acc: torch.Tensor, other: torch.Tensor) -> None:
    if op in (ReduceOp.SUM, ReduceOp.AVG):
        acc.add_(other)
    elif op == ReduceOp.PRODUCT:
        acc.mul_(other)
    elif op == ReduceOp.MIN:
        torch.minimum(acc, other, out=acc)
    elif op == ReduceOp.MAX:
        torch.maximum(acc, other, out=acc)
    elif op == ReduceOp.BAND:
        acc.bitwise_and_(other)
    elif op == ReduceOp.BOR:
        acc.bitwise_or_(other)
    elif op == ReduceOp.BXOR:
        acc.bitwise_xor_(other)
    else:
        raise ValueError(f"unsupported ReduceOp {op}")


def _finish_avg(op: ReduceOp, t: torch.Tensor, world: int) -> None:
    if op == ReduceOp.AVG:
        if t.is_floating_point() or t.is_complex():
            t.div_(world)
        else:
            t.copy_(torch.div(t, world, rounding_mode="floor"))


class CpuBackend:
    name = "cpu"

    def __init__(self, store, rank: int, world_size: int, host: str = "127.0.0.1", prefix: str = "mesh",
                 timeout: float = 1800.0):
        self.rank, self.world = rank, world_size
        self._lib = host_lib()
        self._mesh = None
        if world_size > 1:
            self._mesh = self._lib.dcp_mesh_create(store.client_handle, rank, world_size, host.encode(), prefix.encode(), float(timeout))
            if not self._mesh:
                raise RuntimeError("CpuBackend: could not establish the socket mesh")

    # ---- raw transport ----------------------------------------------------------------------------
    @staticmethod
    def _check(rc: int, what: str) -> None:
        if rc != 0:
            reason = {-1: "connection lost", -2: "size/tag mismatch with peer", -3: "timeout", -4: "bad peer rank"}.get(rc, str(rc))
            raise RuntimeError(f"CpuBackend.{what} failed: {reason}")

    def _send(self, t: torch.Tensor, dst: int, tag: int = 0) -> None:
        self._check(self._lib.dcp_mesh_send(self._mesh, dst, C.c_void_p(t.data_ptr()), t.numel() * t.element_size(), tag), "send")

    def _recv(self, t: torch.Tensor, src: int, tag: int = 0) -> None:
        self._check(self._lib.dcp_mesh_recv(self._mesh, src, C.c_void_p(t.data_ptr()), t.numel() * t.element_size(), tag), "recv")

    def _sendrecv(self, s: torch.Tensor, dst: int, r: torch.Tensor, src: int, tag: int = 0) -> None:
        self._check(self._lib.dcp_mesh_sendrecv(self._mesh, dst, C.c_void_p(s.data_ptr()), s.numel() * s.element_size(),
                                                src, C.c_void_p(r.data_ptr()), r.numel() * r.element_size(), tag), "sendrecv")

    @staticmethod
    def _flat(t: torch.Tensor):
        """Contiguous 1-D alias of ``t`` (or a copy if ``t`` is not contiguous) + whether to copy back."""
        t = t.detach()
        if t.is_contiguous():
            return t.view(-1), False
        return t.contiguous().view(-1), True

    # ---- point to point ---------------------------------------------------------------------------
    def send(self, tensor: torch.Tensor, dst: int, tag: int = 0) -> None:
        self._drain()
        flat, _ = self._flat(tensor)
        self._send(flat, dst, tag)

    def recv(self, tensor: torch.Tensor, src: int, tag: int = 0) -> None:
        self._drain()
        flat, back = self._flat(tensor)
        self._recv(flat, src, tag)
        if back:
            tensor.detach().copy_(flat.view(tensor.shape))

    # Non-blocking point to point: one worker thread owns the sending direction of the sockets and one the receiving direction
    # (the native calls release the GIL; the two directions of a socket are independent byte streams), so the usual exchange -
    # every rank posts irecv, then isend, then waits - makes progress on both.  Operations of one kind complete in the order
    # they were posted; blocking calls and collectives first wait for everything still in flight.
    def _worker(self, kind: str):
        from concurrent.futures import ThreadPoolExecutor
        if getattr(self, "_workers", None) is None:
            self._workers, self._pending = {}, []
        if kind not in self._workers:
            self._workers[kind] = ThreadPoolExecutor(max_workers=1, thread_name_prefix=f"dcp-{kind}")
        return self._workers[kind]

    def _drain(self) -> None:
        pend = getattr(self, "_pending", None)
        if pend:
            self._pending = []
            for f in pend:
                f.result()

    def isend(self, tensor: torch.Tensor, dst: int, tag: int = 0) -> Work:
        flat, _ = self._flat(tensor)
        fut = self._worker("tx").submit(self._send, flat, dst, tag)
        self._pending.append(fut)
        return Work(done=False, finalize=fut.result, poll=fut.done, keep=(tensor, flat))

    def irecv(self, tensor: torch.Tensor, src: int, tag: int = 0) -> Work:
        flat, back = self._flat(tensor)

        def job():
            self._recv(flat, src, tag)
            if back:
                tensor.detach().copy_(flat.view(tensor.shape))
        fut = self._worker("rx").submit(job)
        self._pending.append(fut)
        return Work(done=False, finalize=fut.result, poll=fut.done, keep=(tensor, flat))

    # ---- collectives ------------------------------------------------------------------------------
    def barrier(self) -> None:
        self._drain()
        if self.world == 1:
            return
        token = torch.zeros(1, dtype=torch.int32)
        self.all_reduce(token, ReduceOp.SUM)

    def broadcast(self, tensor: torch.Tensor, src: int) -> None:
        self._drain()
        if self.world == 1:
            return
        flat, back = self._flat(tensor)
        # binomial tree rooted at src
        vrank = (self.rank - src) % self.world
        mask = 1
        while mask < self.world:
            if vrank & mask:
                self._recv(flat, (vrank - mask + src) % self.world, tag=11)
                break
            mask <<= 1
        mask >>= 1
        while mask > 0:
            if vrank + mask < self.world:
                self._send(flat, (vrank + mask + src) % self.world, tag=11)
            mask >>= 1
        if back:
            tensor.detach().copy_(flat.view(tensor.shape))

    def all_reduce(self, tensor: torch.Tensor, op: ReduceOp = ReduceOp.SUM) -> None:
        self._drain()
        if self.world == 1:
            return
        flat, back = self._flat(tensor)
        n, W, r = flat.numel(), self.world, self.rank
        if n * flat.element_size() <= _SMALL_BYTES or n < W:
            # exchange full copies, reduce in rank order (identical on every rank)
            mine = flat.clone()
            parts: List[Optional[torch.Tensor]] = [None] * W
            parts[r] = mine
            for step in range(1, W):
                dst, src = (r + step) % W, (r - step) % W
                buf = torch.empty_like(mine)
                self._sendrecv(mine, dst, buf, src, tag=21)
                parts[src] = buf
            flat.copy_(parts[0])
            for p in parts[1:]:
                _apply(op, flat, p)
        else:
            bounds = [(n * i) // W for i in range(W + 1)]
            chunk = lambda i: flat[bounds[i % W]:bounds[i % W + 1]]
            tmp = torch.empty(max(bounds[i + 1] - bounds[i] for i in range(W)), dtype=flat.dtype)
            right, left = (r + 1) % W, (r - 1) % W
            for step in range(W - 1):  # reduce-scatter
                s, d = chunk(r - step), chunk(r - step - 1)
                rbuf = tmp[: d.numel()]
                self._sendrecv(s, right, rbuf, left, tag=22)
                # keep operand order canonical (lower rank's contribution first) is not needed for
                # cross-rank agreement: each chunk is finished on exactly one rank and then copied.
                _apply(op, d, rbuf)
            for step in range(W - 1):  # all-gather
                s, d = chunk(r + 1 - step), chunk(r - step)
                self._sendrecv(s, right, d, left, tag=23)
        _finish_avg(op, flat, W)
        if back:
            tensor.detach().copy_(flat.view(tensor.shape))

    def reduce(self, tensor: torch.Tensor, dst: int, op: ReduceOp = ReduceOp.SUM) -> None:
        self._drain()
        if self.world == 1:
            return
        flat, back = self._flat(tensor)
        if self.rank == dst:
            acc = [None] * self.world
            acc[dst] = flat.clone()
            for src in range(self.world):
                if src != dst:
                    acc[src] = torch.empty_like(flat)
                    self._recv(acc[src], src, tag=31)
            flat.copy_(acc[0])
            for p in acc[1:]:
                _apply(op, flat, p)
            _finish_avg(op, flat, self.world)
            if back:
                tensor.detach().copy_(flat.view(tensor.shape))
        else:
            self._send(flat, dst, tag=31)

    def all_gather(self, outputs: List[torch.Tensor], tensor: torch.Tensor) -> None:
        self._drain()
        flat, _ = self._flat(tensor)
        W, r = self.world, self.rank
        outputs[r].detach().copy_(tensor.detach())
        for step in range(1, W):
            dst, src = (r + step) % W, (r - step) % W
            out = outputs[src].detach()
            buf = out.view(-1) if out.is_contiguous() else torch.empty(out.numel(), dtype=out.dtype)
            self._sendrecv(flat, dst, buf, src, tag=41)
            if not out.is_contiguous():
                out.copy_(buf.view(out.shape))

    def all_gather_into_tensor(self, output: torch.Tensor, tensor: torch.Tensor) -> None:
        self._drain()
        out = output.detach()
        assert out.is_contiguous() and out.numel() == tensor.numel() * self.world
        self.all_gather(list(out.view(-1).chunk(self.world)), tensor.detach().reshape(-1))

    def gather(self, tensor: torch.Tensor, gather_list: Optional[List[torch.Tensor]], dst: int) -> None:
        self._drain()
        flat, _ = self._flat(tensor)
        if self.rank == dst:
            for src in range(self.world):
                out = gather_list[src].detach()
                if src == dst:
                    out.copy_(tensor.detach())
                else:
                    buf = out.view(-1) if out.is_contiguous() else torch.empty(out.numel(), dtype=out.dtype)
                    self._recv(buf, src, tag=51)
                    if not out.is_contiguous():
                        out.copy_(buf.view(out.shape))
        else:
            self._send(flat, dst, tag=51)

    def scatter(self, tensor: torch.Tensor, scatter_list: Optional[List[torch.Tensor]], src: int) -> None:
        self._drain()
        if self.rank == src:
            for dst in range(self.world):
                if dst == src:
                    tensor.detach().copy_(scatter_list[dst].detach())
                else:
                    flat, _ = self._flat(scatter_list[dst])
                    self._send(flat, dst, tag=61)
        else:
            self.recv(tensor, src, tag=61)

    def reduce_scatter(self, output: torch.Tensor, input_list: List[torch.Tensor], op: ReduceOp = ReduceOp.SUM) -> None:
        self._drain()
        W, r = self.world, self.rank
        acc: List[Optional[torch.Tensor]] = [None] * W
        acc[r] = input_list[r].detach().contiguous().view(-1).clone()
        for step in range(1, W):
            dst, src = (r + step) % W, (r - step) % W
            s, _ = self._flat(input_list[dst])
            buf = torch.empty_like(acc[r])
            self._sendrecv(s, dst, buf, src, tag=71)
            acc[src] = buf
        res = acc[0]
        for p in acc[1:]:
            _apply(op, res, p)
        _finish_avg(op, res, W)
        output.detach().copy_(res.view(output.shape))

    def reduce_scatter_tensor(self, output: torch.Tensor, input: torch.Tensor, op: ReduceOp = ReduceOp.SUM) -> None:
        self._drain()
        inp = input.detach().contiguous().view(-1)
        self.reduce_scatter(output, list(inp.chunk(self.world)), op)

    def close(self) -> None:
        for ex in (getattr(self, "_workers", None) or {}).values():
            ex.shutdown(wait=True)
        self._workers = None
        if self._mesh:
            self._lib.dcp_mesh_destroy(self._mesh)
            self._mesh = None
