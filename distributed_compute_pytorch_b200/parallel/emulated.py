"""An emulated multi-rank world on ONE GPU - the single-GPU test tier of the NVLink collectives.

The collective kernels of csrc/coll wait on each other across ranks (signal words, parity halves).  Kernels that wait on
one another must not be issued as separate launches on one GPU - nothing guarantees they are resident together - so
"N processes on one device" is not a way to test them.  Instead every kernel body can also run as ONE cooperative launch
in which ``blockIdx.y`` selects the rank (csrc/coll/coll_launch.cuh): all ranks' CTAs are co-resident by construction, and
the code that runs is the production body - same signal protocol, same double-buffering, same reductions - with the W
symmetric heaps living on the same device and "peer" pointers being plain device pointers.  (No NVLS in this mode: the
multimem paths are covered by the multi-GPU tests.)

    world = EmulatedWorld(4)
    for r, be in enumerate(world.ranks):       # every rank issues the collective, as it would in its own process
        be.all_reduce(t[r])                    # ... which only QUEUES it
    world.flush()                              # one cooperative kernel per queued collective, all ranks inside

``world.ranks[r]`` is a :class:`~.p2p_backend.P2PBackend` in every other respect (``empty_symmetric``, ``handle``, ...), so
the C entry points that take a heap handle (``dcp_fused_rs_opt_ag`` etc.) work on it too.  What the reference does with
``dist.all_reduce`` over gloo (ref:main.py:65,90,91) is what these kernels do on a B200; this module is how their protocol
logic gets exercised when only one GPU is present.
"""
from __future__ import annotations

import ctypes as C
from typing import List

import torch

from .._native import cuda_lib, host_lib
from .p2p_backend import P2PBackend


def _declare(lib) -> None:
    if getattr(lib, "_dcp_emu_declared", False):
        return
    vp, i, sz, d = C.c_void_p, C.c_int, C.c_size_t, C.c_double
    lib.dcp_emu_create.restype, lib.dcp_emu_create.argtypes = vp, [i, i, sz, d, sz, sz]
    lib.dcp_emu_rank.restype, lib.dcp_emu_rank.argtypes = vp, [vp, i]
    lib.dcp_emu_pending.restype, lib.dcp_emu_pending.argtypes = i, [vp]
    lib.dcp_emu_flush.restype, lib.dcp_emu_flush.argtypes = i, [vp, vp]
    lib.dcp_emu_destroy.restype, lib.dcp_emu_destroy.argtypes = None, [vp]
    lib._dcp_emu_declared = True


class EmulatedRank(P2PBackend):
    """One rank of an :class:`EmulatedWorld`.  Collectives are queued; nothing runs before ``world.flush()``."""

    def __init__(self, world: "EmulatedWorld", rank: int):      # noqa: D401 - deliberately does not call P2PBackend.__init__
        self.emu = world
        self.rank, self.world, self.device = rank, world.size, world.device
        self.store = None
        self.lib, self.hlib = world.lib, host_lib()
        self.timeout = world.timeout
        self._torch_dev = torch.device("cuda", self.device)
        self._h = self.lib.dcp_emu_rank(world.group, rank)
        self.mode, self.nvls = 1, False
        self.heap_base = self.lib.dcp_symm_local_ptr(self._h)
        self.heap_bytes = self.lib.dcp_symm_bytes(self._h)
        self._user_off = self.lib.dcp_symm_user_offset(self._h)
        self._free = [(self._user_off, self.heap_bytes - self._user_off)]
        self._live = {}
        self.oneshot_max = min(max(8 * 1024, (512 << 10) // max(1, (self.world - 1) ** 2)), self.lib.dcp_coll_oneshot_max_bytes(self._h))
        self.nvls_min = 1 << 62
        self.comm_stream = torch.cuda.current_stream(self._torch_dev)
        self._algo_override = 0
        self.launches = 0
        self.default_max_ctas = 4        # world x CTAs must be co-resident in the cooperative launch

    def barrier(self) -> None:
        """Queues the device barrier (no host synchronisation: the kernel has not been launched yet)."""
        if self.world > 1:
            from .p2p_backend import CH_USER
            self._check(self.lib.dcp_barrier_launch(self._h, CH_USER, self._stream()), "barrier")

    def close(self) -> None:       # the world owns the heaps
        self._h = None


class EmulatedWorld:
    """``size`` ranks on ``device``; ``ranks[r]`` queues collectives, ``flush()`` runs them (all ranks in one launch each)."""

    def __init__(self, size: int, device: int = 0, heap_mb: int = 192, staging_mb: int = 64, p2p_slot_mb: int = 2, timeout: float = 20.0):
        self.lib = cuda_lib()
        _declare(self.lib)
        self.size, self.device, self.timeout = int(size), int(device), float(timeout)
        with torch.cuda.device(self.device):
            torch.zeros(1, device=torch.device("cuda", self.device))
            self.group = self.lib.dcp_emu_create(self.size, self.device, heap_mb << 20, self.timeout, p2p_slot_mb << 20, staging_mb << 20)
        if not self.group:
            raise RuntimeError(f"could not create the emulated world: {(self.lib.dcp_last_error() or b'').decode()}")
        self.ranks: List[EmulatedRank] = [EmulatedRank(self, r) for r in range(self.size)]

    def pending(self) -> int:
        return self.lib.dcp_emu_pending(self.group)

    def flush(self, stream=None) -> None:
        dev = torch.device("cuda", self.device)
        st = (stream or torch.cuda.current_stream(dev)).cuda_stream
        with torch.cuda.device(dev):
            rc = self.lib.dcp_emu_flush(self.group, st)
        if rc != 0:
            raise RuntimeError(f"emulated world: {(self.lib.dcp_last_error() or b'').decode()}")

    def check_health(self) -> None:
        """Raise if any rank's bounded wait expired (call after a synchronize)."""
        for be in self.ranks:
            if self.lib.dcp_symm_error_flag(be.handle):
                raise RuntimeError(f"emulated rank {be.rank}: a wait timed out (collective mismatch)")

    def close(self) -> None:
        if self.group:
            for be in self.ranks:
                be.close()
            with torch.cuda.device(self.device):
                torch.cuda.synchronize()
                self.lib.dcp_emu_destroy(self.group)
            self.group = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()
