"""torch.distributed-shaped process-group API.

Call-compatible with what the reference uses - ``init_process_group(backend, rank=, world_size=)``,
``destroy_process_group()``, ``all_reduce(t, op=ReduceOp.SUM)`` (ref:main.py:50,53,65,90,91) - plus
the rest of the primitive set (send/recv, broadcast, reduce, all_gather, gather, scatter,
reduce_scatter, barrier).  CUDA tensors are routed to the hand-written NVLink P2P backend
(``parallel/p2p_backend.py``), CPU tensors to the socket mesh (``cpu_backend.py``); NCCL and gloo are
never called.
"""
from __future__ import annotations

import os
from datetime import timedelta
from typing import List, Optional

import torch

from .store import DEFAULT_TIMEOUT_S, PrefixStore, TCPStore, rendezvous
from .types import ReduceOp, Work

_CPU_NAMES = {"gloo", "cpu", "mesh"}
_GPU_NAMES = {"nccl", "b200", "nvlink", "p2p", "cuda"}


class ProcessGroup:
    def __init__(self, rank, world_size, store, backend_name, cpu, cuda, owns_store):
        self.rank_, self.size_, self.store, self.backend_name = rank, world_size, store, backend_name
        self.cpu, self.cuda, self.owns_store = cpu, cuda, owns_store
        self.pstore = None

    def rank(self) -> int: return self.rank_
    def size(self) -> int: return self.size_
    def name(self) -> str: return "world"


class _GroupMember:
    WORLD: Optional[ProcessGroup] = None


group = _GroupMember
_pg_counter = 0


def is_available() -> bool:
    return True


def is_initialized() -> bool:
    return _GroupMember.WORLD is not None


def _world(g=None) -> ProcessGroup:
    if g is not None and g is not _GroupMember.WORLD:
        raise NotImplementedError("only the default (world) group is supported")
    pg = _GroupMember.WORLD
    if pg is None:
        raise RuntimeError("Default process group has not been initialized, please make sure to call init_process_group.")
    return pg


def _timeout_s(timeout) -> float:
    if timeout is None:
        return DEFAULT_TIMEOUT_S
    if isinstance(timeout, timedelta):
        return timeout.total_seconds()
    return float(timeout)


def init_process_group(backend: Optional[str] = None, init_method: Optional[str] = None, timeout=None,
                       world_size: int = -1, rank: int = -1, store=None, group_name: str = "",
                       pg_options=None, device_id=None) -> None:
    """Join the job.  Same signature as the library call the reference makes at ref:main.py:50.

    ``backend``: ``"gloo"``/``"cpu"`` -> CPU mesh only; ``"nccl"``/``"b200"`` -> NVLink P2P backend for
    CUDA tensors plus the CPU mesh for CPU tensors; ``None`` -> GPU backend iff CUDA is available.
    The reference hard-codes ``"gloo"`` even on GPUs; with CUDA present we still bring up the P2P
    backend lazily-free (eagerly) so CUDA tensors never bounce through the host.
    """
    global _pg_counter
    if is_initialized():
        raise RuntimeError("trying to initialize the default process group twice!")
    tmo = _timeout_s(timeout)
    owns_store = store is None
    if store is None:
        store, rank, world_size = rendezvous(init_method, rank, world_size, tmo)
    elif rank < 0 or world_size < 0:
        raise ValueError("rank and world_size are required when a store is passed")
    name = (backend or "").lower()
    if name and name not in _CPU_NAMES | _GPU_NAMES:
        raise ValueError(f"unknown backend {backend!r}")
    use_cuda = torch.cuda.is_available() and name not in ("cpu", "mesh") and os.environ.get("DCP_DISABLE_CUDA") != "1"
    if name in _GPU_NAMES and not torch.cuda.is_available():
        raise RuntimeError(f"backend {backend!r} needs CUDA, which is not available")
    pstore = PrefixStore(f"pg{_pg_counter}", store)
    _pg_counter += 1

    from .cpu_backend import CpuBackend
    host = os.environ.get("DCP_HOST_ADDR", "127.0.0.1" if getattr(store, "host", "127.0.0.1") in ("127.0.0.1", "localhost") else _my_addr())
    cpu = CpuBackend(pstore, rank, world_size, host=host, timeout=tmo)
    cuda = None
    if use_cuda:
        from .parallel.p2p_backend import P2PBackend
        if device_id is None:
            if "LOCAL_RANK" in os.environ:
                dev = int(os.environ["LOCAL_RANK"])
            else:
                dev = rank % torch.cuda.device_count()
        else:
            dev = torch.device(device_id).index if not isinstance(device_id, int) else device_id
        cuda = P2PBackend(pstore, rank, world_size, dev, timeout=tmo)
    _GroupMember.WORLD = ProcessGroup(rank, world_size, store, name or ("b200" if use_cuda else "cpu"), cpu, cuda, owns_store)
    _GroupMember.WORLD.pstore = pstore
    pstore.barrier("init", rank, world_size, tmo)


def _my_addr() -> str:
    import socket
    try:
        return socket.gethostbyname(socket.gethostname())
    except OSError:
        return "127.0.0.1"


def destroy_process_group(group=None) -> None:
    """Tear down (ref:main.py:52-53): unmap peers, close sockets, stop the store server last."""
    pg = _GroupMember.WORLD
    if pg is None:
        return
    try:
        if pg.cuda is not None:
            torch.cuda.synchronize()
        pg.pstore.barrier("destroy", pg.rank_, pg.size_, 60.0)
    except Exception:
        pass
    if pg.cuda is not None:
        pg.cuda.close()
    pg.cpu.close()
    _GroupMember.WORLD = None
    if pg.owns_store:
        if pg.rank_ == 0:
            # let clients finish their last request before the server disappears
            import time
            time.sleep(0.05)
        pg.store.close()


def get_rank(group=None) -> int:
    return _world(group).rank_


def get_world_size(group=None) -> int:
    return _world(group).size_


def get_backend(group=None) -> str:
    return _world(group).backend_name


def _backend_for(t: torch.Tensor, pg: ProcessGroup):
    if t.is_cuda:
        if pg.cuda is None:
            raise RuntimeError("CUDA tensor passed to a process group initialised without the GPU backend")
        return pg.cuda
    return pg.cpu


def _done(async_op: bool, be=None):
    if not async_op:
        return None
    if be is not None and hasattr(be, "work"):
        return be.work()
    return Work()


# ---- point to point -------------------------------------------------------------------------------
def send(tensor: torch.Tensor, dst: Optional[int] = None, group=None, tag: int = 0, group_dst: Optional[int] = None) -> None:
    pg = _world(group)
    dst = dst if dst is not None else group_dst
    if dst == pg.rank_:
        raise ValueError("Invalid destination rank: destination rank should not be the same as the rank of the current process.")
    _backend_for(tensor, pg).send(tensor, dst, tag)


def recv(tensor: torch.Tensor, src: Optional[int] = None, group=None, tag: int = 0, group_src: Optional[int] = None) -> int:
    pg = _world(group)
    src = src if src is not None else group_src
    if src is None:
        raise NotImplementedError("recv from any source is not supported; pass src")
    _backend_for(tensor, pg).recv(tensor, src, tag)
    return src


def isend(tensor, dst=None, group=None, tag=0, group_dst=None) -> Work:
    """Non-blocking send: returns at once; ``wait()`` on the handle completes it (CPU: joins the sender thread; CUDA: orders the
    caller's stream behind the transfer).  The tensor must not be modified before that."""
    pg = _world(group)
    dst = dst if dst is not None else group_dst
    if dst == pg.rank_:
        raise ValueError("Invalid destination rank: destination rank should not be the same as the rank of the current process.")
    return _backend_for(tensor, pg).isend(tensor, dst, tag)


def irecv(tensor, src=None, group=None, tag=0, group_src=None) -> Work:
    """Non-blocking receive: the tensor holds the data once ``wait()`` on the handle has returned."""
    pg = _world(group)
    src = src if src is not None else group_src
    if src is None:
        raise NotImplementedError("irecv from any source is not supported; pass src")
    return _backend_for(tensor, pg).irecv(tensor, src, tag)


# ---- collectives ----------------------------------------------------------------------------------
def broadcast(tensor: torch.Tensor, src: Optional[int] = None, group=None, async_op: bool = False, group_src: Optional[int] = None):
    pg = _world(group)
    src = src if src is not None else (group_src if group_src is not None else 0)
    be = _backend_for(tensor, pg)
    be.broadcast(tensor, src)
    return _done(async_op, be)


def all_reduce(tensor: torch.Tensor, op: ReduceOp = ReduceOp.SUM, group=None, async_op: bool = False):
    """In-place all-reduce.  Accepts what the reference passes: a 0-dim fp32 tensor that requires grad
    (ref:main.py:65), a 0-dim fp32 (ref:main.py:90) and a 0-dim int64 (ref:main.py:91)."""
    pg = _world(group)
    be = _backend_for(tensor, pg)
    be.all_reduce(tensor, ReduceOp(int(op)))
    return _done(async_op, be)


def reduce(tensor: torch.Tensor, dst: Optional[int] = None, op: ReduceOp = ReduceOp.SUM, group=None, async_op: bool = False,
           group_dst: Optional[int] = None):
    pg = _world(group)
    dst = dst if dst is not None else (group_dst if group_dst is not None else 0)
    be = _backend_for(tensor, pg)
    be.reduce(tensor, dst, ReduceOp(int(op)))
    return _done(async_op, be)


def all_gather(tensor_list: List[torch.Tensor], tensor: torch.Tensor, group=None, async_op: bool = False):
    pg = _world(group)
    if len(tensor_list) != pg.size_:
        raise ValueError("tensor_list must have world_size entries")
    be = _backend_for(tensor, pg)
    be.all_gather(tensor_list, tensor)
    return _done(async_op, be)


def all_gather_into_tensor(output_tensor: torch.Tensor, input_tensor: torch.Tensor, group=None, async_op: bool = False):
    pg = _world(group)
    be = _backend_for(input_tensor, pg)
    be.all_gather_into_tensor(output_tensor, input_tensor)
    return _done(async_op, be)


def gather(tensor: torch.Tensor, gather_list: Optional[List[torch.Tensor]] = None, dst: Optional[int] = None, group=None,
           async_op: bool = False, group_dst: Optional[int] = None):
    pg = _world(group)
    dst = dst if dst is not None else (group_dst if group_dst is not None else 0)
    if pg.rank_ == dst and (gather_list is None or len(gather_list) != pg.size_):
        raise ValueError("gather_list with world_size entries is required on the destination rank")
    be = _backend_for(tensor, pg)
    be.gather(tensor, gather_list, dst)
    return _done(async_op, be)


def scatter(tensor: torch.Tensor, scatter_list: Optional[List[torch.Tensor]] = None, src: Optional[int] = None, group=None,
            async_op: bool = False, group_src: Optional[int] = None):
    pg = _world(group)
    src = src if src is not None else (group_src if group_src is not None else 0)
    if pg.rank_ == src and (scatter_list is None or len(scatter_list) != pg.size_):
        raise ValueError("scatter_list with world_size entries is required on the source rank")
    be = _backend_for(tensor, pg)
    be.scatter(tensor, scatter_list, src)
    return _done(async_op, be)


def reduce_scatter(output: torch.Tensor, input_list: List[torch.Tensor], op: ReduceOp = ReduceOp.SUM, group=None,
                   async_op: bool = False):
    pg = _world(group)
    if len(input_list) != pg.size_:
        raise ValueError("input_list must have world_size entries")
    be = _backend_for(output, pg)
    be.reduce_scatter(output, input_list, ReduceOp(int(op)))
    return _done(async_op, be)


def reduce_scatter_tensor(output: torch.Tensor, input: torch.Tensor, op: ReduceOp = ReduceOp.SUM, group=None,
                          async_op: bool = False):
    pg = _world(group)
    be = _backend_for(output, pg)
    be.reduce_scatter_tensor(output, input, ReduceOp(int(op)))
    return _done(async_op, be)


def barrier(group=None, async_op: bool = False, device_ids=None):
    pg = _world(group)
    if pg.cuda is not None:
        pg.cuda.barrier()
    else:
        pg.cpu.barrier()
    return _done(async_op)


def broadcast_object_list(object_list: list, src: int = 0, group=None) -> None:
    """Pickle-broadcast small Python objects over the store (host side; not a hot path)."""
    pg = _world(group)
    seq = getattr(pg, "_bol_seq", 0)
    pg._bol_seq = seq + 1
    key = f"bol/{seq}"
    if pg.rank_ == src:
        pg.store.set_obj(key, list(object_list))
    else:
        object_list[:] = pg.store.get_obj(key)
