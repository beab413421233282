"""NVLink P2P backend: the CUDA side of the process group.

Brings up the symmetric heap (csrc/symm/symm.cu: cuMem VMM + POSIX-FD exchange over Unix sockets,
NVLS multicast when the driver allows it, cudaIpc as fallback), then routes every collective on a
CUDA tensor to the hand-written sm_100a kernels in csrc/coll/coll.cu.  NCCL is never called.
This is what stands behind ``dist.all_reduce`` at ref:main.py:65,90,91 and behind the DDP gradient
buckets (ref:main.py:62,122) when the tensors live on a B200.

Algorithm choice for all_reduce (per message size, override with DCP_ALLREDUCE_ALGO=1|2|3):
  <= DCP_ONESHOT_BYTES (default 512 KiB/(W-1)^2) one-shot push   (latency floor: 1 NVLink hop + 1 signal)
  larger, multicast mapped, f32/bf16/f16   NVLS            (multimem.ld_reduce + multimem.st)
  otherwise                                two-shot P2P    (pull-reduce own slice, push to all)
Tensors inside the symmetric heap (``empty_symmetric``; DDP buckets) are reduced in place with zero
copies; other tensors bounce through a double-buffered staging area inside the same kernel.
"""
from __future__ import annotations

import ctypes as C
import os
import threading
from typing import Dict, List, Optional, Tuple

import torch

from .._native import cuda_lib, host_lib
from ..types import ReduceOp, Work

_DTYPE = {torch.float32: 0, torch.bfloat16: 1, torch.float16: 2, torch.int32: 3, torch.int64: 4, torch.float64: 5,
          torch.uint8: 6, torch.bool: 6, torch.int8: 6}

CH_USER, CH_DDP, CH_BN, CH_OPT, CH_TAIL = 0, 1, 2, 3, 4  # signal-pad channels (independent epoch streams)


class _CudaBuf:
    """Exposes a raw device pointer to torch through __cuda_array_interface__ (zero copy)."""

    def __init__(self, ptr: int, nbytes: int, owner):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3, "strides": None}
        self._owner = owner


class P2PBackend:
    name = "b200"

    def __init__(self, store, rank: int, world_size: int, device: int, timeout: float = 1800.0):
        self.rank, self.world, self.device = rank, world_size, int(device)
        self.store = store
        self.lib = cuda_lib()
        self.hlib = host_lib()
        self.timeout = float(os.environ.get("DCP_COLL_TIMEOUT_S", min(timeout, 120.0)))
        heap_mb = int(os.environ.get("DCP_HEAP_MB", "2048"))
        staging_mb = int(os.environ.get("DCP_STAGING_MB", "288"))
        p2p_slot = int(os.environ.get("DCP_P2P_SLOT_MB", "16")) << 20   # two 8 MiB halves per (src, dst) pair
        mode_env = os.environ.get("DCP_SYMM", "vmm")
        self._torch_dev = torch.device("cuda", self.device)
        with torch.cuda.device(self._torch_dev):
            torch.cuda.init()
            torch.zeros(1, device=self._torch_dev)  # make sure the primary context exists
            self._h = None
            self.mode = None
            for mode in ([0, 1] if mode_env == "vmm" else [1]):
                h = self.lib.dcp_symm_create(rank, world_size, self.device, heap_mb << 20, mode)
                ok = 1 if h else 0
                if self._agree(f"symm_create_{mode}", ok):
                    self._h, self.mode = h, mode
                    break
                if h:
                    self.lib.dcp_symm_destroy(h)
            if self._h is None:
                raise RuntimeError(f"could not create the symmetric heap: {self._err()}")
            nvls_ok = self._exchange_vmm() if self.mode == 0 else self._exchange_ipc()
            if os.environ.get("DCP_DISABLE_NVLS") == "1":
                nvls_ok = False
            if self.lib.dcp_symm_finalize(self._h, self.timeout, p2p_slot, staging_mb << 20, 1 if nvls_ok else 0) != 0:
                raise RuntimeError(f"symmetric heap finalize failed: {self._err()}")
        self.nvls = bool(self.lib.dcp_symm_has_nvls(self._h))
        self.heap_base = self.lib.dcp_symm_local_ptr(self._h)
        self.heap_bytes = self.lib.dcp_symm_bytes(self._h)
        self._user_off = self.lib.dcp_symm_user_offset(self._h)
        self._free: List[Tuple[int, int]] = [(self._user_off, self.heap_bytes - self._user_off)]
        self._live: Dict[int, Tuple[int, int]] = {}
        # one-shot pushes (W-1) full copies per rank, two-shot moves 2(W-1)/W: the crossover shrinks with W
        # measured crossovers (profiles/coll_sweep_*): ~512 KiB at W=2, ~10 KiB at W=8
        default_oneshot = max(8 * 1024, (512 << 10) // max(1, (world_size - 1) ** 2))
        self.oneshot_max = min(int(os.environ.get("DCP_ONESHOT_BYTES", str(default_oneshot))), self.lib.dcp_coll_oneshot_max_bytes(self._h))
        self.nvls_min = int(os.environ.get("DCP_NVLS_MIN_BYTES", str(256 * 1024)))
        self.comm_stream = torch.cuda.Stream(device=self._torch_dev, priority=-1)
        self._algo_override = int(os.environ.get("DCP_ALLREDUCE_ALGO", "0"))
        self.launches = 0
        self.default_max_ctas = 0        # 0: the native side sizes the grid from the message
        store.barrier("p2p_ready", rank, world_size, timeout)

    # ------------------------------------------------------------------ bring-up -----------------
    def _err(self) -> str:
        return (self.lib.dcp_last_error() or b"").decode()

    def _agree(self, key: str, ok: int) -> bool:
        """True iff every rank reports ok."""
        self.store.set(f"{key}/{self.rank}", b"1" if ok else b"0")
        return all(self.store.get(f"{key}/{r}") == b"1" for r in range(self.world))

    def _exchange_vmm(self) -> bool:
        W, r, lib, hl = self.world, self.rank, self.lib, self.hlib
        if W == 1:
            return False
        if r == 0:
            import uuid
            self.store.set("jobid", uuid.uuid4().hex[:12])
        job = self.store.get("jobid").decode()
        name = lambda k: f"dcp-{job}-{k}".encode()
        lfd = hl.dcp_fd_listen(name(r))
        if lfd < 0:
            raise RuntimeError("cannot open the FD-passing socket")
        self.store.barrier("fd_listen", r, W, self.timeout)
        want_mc = bool(lib.dcp_symm_multicast_supported(self.device)) and os.environ.get("DCP_DISABLE_NVLS") != "1"
        my_fd = lib.dcp_symm_export_fd(self._h)
        if my_fd < 0:
            raise RuntimeError(f"export of the heap handle failed: {self._err()}")
        mc_fd = -1
        if r == 0 and want_mc:
            mc_fd = lib.dcp_symm_mc_create(self._h)
        got: Dict[int, List[int]] = {}
        errs: List[str] = []

        def receiver():
            for _ in range(W - 1):
                fds = (C.c_int * 4)()
                n = C.c_int(0)
                src = hl.dcp_fd_recv(lfd, fds, 4, C.byref(n), self.timeout)
                if src < 0:
                    errs.append(f"fd recv failed ({src})")
                    return
                got[src] = list(fds[: n.value])

        t = threading.Thread(target=receiver, daemon=True)
        t.start()
        for p in range(W):
            if p == r:
                continue
            fds = [my_fd] + ([mc_fd] if (r == 0 and mc_fd >= 0) else [])
            arr = (C.c_int * len(fds))(*fds)
            rc = hl.dcp_fd_send(name(p), r, arr, len(fds), self.timeout)
            if rc != 0:
                raise RuntimeError(f"sending heap handle to rank {p} failed ({rc})")
        t.join()
        hl.dcp_fd_close(lfd)
        if errs:
            raise RuntimeError(errs[0])
        for p, fds in got.items():
            if lib.dcp_symm_import_fd(self._h, p, fds[0]) != 0:
                raise RuntimeError(f"mapping rank {p}'s heap failed: {self._err()}")
            os.close(fds[0])
        os.close(my_fd)
        # ---- NVLS multicast: create (rank 0) -> add device (all) -> bind + map (all) ----
        ok = 1
        if r == 0:
            ok = 1 if mc_fd >= 0 else 0
            if mc_fd >= 0:
                os.close(mc_fd)
        else:
            fds0 = got.get(0, [])
            if len(fds0) >= 2:
                ok = 1 if lib.dcp_symm_mc_import(self._h, fds0[1]) == 0 else 0
                os.close(fds0[1])
            else:
                ok = 0
        if not self._agree("mc_have", ok):
            return False
        ok = 1 if lib.dcp_symm_mc_add_device(self._h) == 0 else 0
        if not self._agree("mc_added", ok):
            return False
        ok = 1 if lib.dcp_symm_mc_bind(self._h) == 0 else 0
        return self._agree("mc_bound", ok)

    def _exchange_ipc(self) -> bool:
        W, r, lib = self.world, self.rank, self.lib
        buf = C.create_string_buffer(64)
        if lib.dcp_symm_ipc_handle(self._h, buf) != 0:
            raise RuntimeError(f"cudaIpcGetMemHandle failed: {self._err()}")
        self.store.set(f"ipc/{r}", buf.raw)
        for p in range(W):
            if p == r:
                continue
            hb = C.create_string_buffer(self.store.get(f"ipc/{p}"), 64)
            if lib.dcp_symm_ipc_open(self._h, p, hb) != 0:
                raise RuntimeError(f"cudaIpcOpenMemHandle(rank {p}) failed: {self._err()}")
        return False

    # ------------------------------------------------------------ symmetric allocator ------------
    def empty_symmetric(self, numel: int, dtype: torch.dtype) -> torch.Tensor:
        """Tensor inside the symmetric heap.  Every rank must allocate in the same order and size."""
        es = torch.empty((), dtype=dtype).element_size()
        nbytes = max(256, -(-numel * es // 256) * 256)
        for i, (off, size) in enumerate(self._free):
            if size >= nbytes:
                if size == nbytes:
                    self._free.pop(i)
                else:
                    self._free[i] = (off + nbytes, size - nbytes)
                break
        else:
            raise RuntimeError(f"symmetric heap exhausted ({nbytes} bytes requested); raise DCP_HEAP_MB")
        ptr = self.heap_base + off
        self._live[ptr] = (off, nbytes)
        raw = torch.as_tensor(_CudaBuf(ptr, nbytes, self), device=self._torch_dev)
        return raw[: numel * es].view(dtype)

    def free_symmetric(self, t: torch.Tensor) -> None:
        ent = self._live.pop(t.data_ptr(), None)
        if ent is None:
            return
        self._free.append(ent)
        self._free.sort()
        merged: List[Tuple[int, int]] = []
        for off, size in self._free:
            if merged and merged[-1][0] + merged[-1][1] == off:
                merged[-1] = (merged[-1][0], merged[-1][1] + size)
            else:
                merged.append((off, size))
        self._free = merged

    def is_symmetric(self, t: torch.Tensor) -> bool:
        return self.heap_base <= t.data_ptr() < self.heap_base + self.heap_bytes

    # --------------------------------------------------------------------- helpers ---------------
    def _stream(self, stream=None) -> int:
        return (stream or torch.cuda.current_stream(self._torch_dev)).cuda_stream

    def _check(self, rc: int, what: str) -> None:
        self.launches += 1
        if rc != 0:
            raise RuntimeError(f"{what} failed: {self._err()}")
        if self.lib.dcp_symm_error_flag(self._h):
            raise RuntimeError(f"{what}: a peer did not arrive within {self.timeout}s (collective mismatch or dead rank)")

    def check_health(self) -> None:
        if self.lib.dcp_symm_error_flag(self._h):
            raise RuntimeError(f"a collective timed out after {self.timeout}s (collective mismatch or dead rank)")

    @staticmethod
    def _code(t: torch.Tensor) -> int:
        try:
            return _DTYPE[t.dtype]
        except KeyError:
            raise TypeError(f"dtype {t.dtype} is not supported by the P2P backend")

    def _contig(self, t: torch.Tensor):
        t = t.detach()
        if t.device != self._torch_dev:
            raise RuntimeError(f"tensor on {t.device} but this rank's process group is bound to {self._torch_dev}")
        if t.is_contiguous():
            return t, False
        return t.contiguous(), True

    def work(self) -> Work:
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self._torch_dev))
        return Work(event=ev)

    def pick_allreduce_algo(self, t: torch.Tensor, op: ReduceOp, nbytes: int) -> int:
        if self._algo_override:
            return self._algo_override
        if nbytes <= self.oneshot_max:
            return 1
        # NVLS pays from three ranks up: with two, the pull-reduce / push of the P2P two-shot moves the same bytes at link rate
        # while a multicast access costs a round trip through the switch (measured at 256 MiB, 2 GPUs: 435 us vs 710 us)
        if self.nvls and self.world > 2 and nbytes >= self.nvls_min and t.dtype in (torch.float32, torch.bfloat16, torch.float16) and op in (ReduceOp.SUM, ReduceOp.AVG):
            return 3
        return 2

    # ------------------------------------------------------------------- collectives -------------
    def all_reduce(self, tensor: torch.Tensor, op: ReduceOp = ReduceOp.SUM, stream=None, algo: int = 0, chan: int = CH_USER,
                   max_ctas: int = 0) -> None:
        t, back = self._contig(tensor)
        if t.dtype in (torch.bool, torch.int8):
            raise TypeError("all_reduce on bool/int8 tensors is not supported")
        if self.world == 1:
            return
        nbytes = t.numel() * t.element_size()
        algo = algo or self.pick_allreduce_algo(t, op, nbytes)
        if max_ctas <= 0:
            max_ctas = self.default_max_ctas or (16 if nbytes <= (1 << 20) else (64 if nbytes <= (32 << 20) else 148))
        with torch.cuda.device(self._torch_dev):
            rc = self.lib.dcp_all_reduce(self._h, t.data_ptr(), t.numel(), self._code(t), int(op), algo, chan, max_ctas, self._stream(stream))
        self._check(rc, "all_reduce")
        if back:
            tensor.detach().copy_(t)

    def all_reduce_async(self, tensor: torch.Tensor, op: ReduceOp = ReduceOp.AVG, max_ctas: Optional[int] = None):
        """Gradient-bucket path: run on the side stream after the producer stream, return an event."""
        cur = torch.cuda.current_stream(self._torch_dev)
        self.comm_stream.wait_stream(cur)
        self.all_reduce(tensor, op, stream=self.comm_stream, chan=CH_DDP, max_ctas=max_ctas or 0)
        ev = torch.cuda.Event()
        ev.record(self.comm_stream)
        return ev

    def broadcast(self, tensor: torch.Tensor, src: int, stream=None, chan: int = CH_USER) -> None:
        t, back = self._contig(tensor)
        if self.world == 1:
            return
        with torch.cuda.device(self._torch_dev):
            rc = self.lib.dcp_broadcast(self._h, t.data_ptr(), t.numel() * t.element_size(), src, chan, self.default_max_ctas, self._stream(stream))
        self._check(rc, "broadcast")
        if back:
            tensor.detach().copy_(t)

    def reduce(self, tensor: torch.Tensor, dst: int, op: ReduceOp = ReduceOp.SUM, stream=None) -> None:
        t, back = self._contig(tensor)
        if self.world == 1:
            return
        out = t if self.rank == dst else t  # non-root output is never written
        with torch.cuda.device(self._torch_dev):
            rc = self.lib.dcp_reduce_scatter(self._h, t.data_ptr(), out.data_ptr(), t.numel(), self._code(t), int(op), dst, CH_USER, self.default_max_ctas, self._stream(stream))
        self._check(rc, "reduce")
        if back and self.rank == dst:
            tensor.detach().copy_(t)

    def all_gather_into_tensor(self, output: torch.Tensor, tensor: torch.Tensor, stream=None) -> None:
        t, _ = self._contig(tensor)
        out = output.detach()
        if not out.is_contiguous() or out.numel() != t.numel() * self.world or out.dtype != t.dtype:
            raise ValueError("all_gather_into_tensor: output must be contiguous with world_size * input.numel() elements")
        if self.world == 1:
            out.view(-1).copy_(t.view(-1))
            return
        with torch.cuda.device(self._torch_dev):
            rc = self.lib.dcp_all_gather(self._h, t.data_ptr(), out.data_ptr(), t.numel() * t.element_size(), -1, CH_USER, self.default_max_ctas, self._stream(stream))
        self._check(rc, "all_gather")

    def all_gather(self, outputs: List[torch.Tensor], tensor: torch.Tensor) -> None:
        flat = torch.empty(self.world * tensor.numel(), dtype=tensor.dtype, device=tensor.device)
        self.all_gather_into_tensor(flat, tensor)
        for r, o in enumerate(outputs):
            o.detach().copy_(flat[r * tensor.numel(): (r + 1) * tensor.numel()].view(o.shape))

    def gather(self, tensor: torch.Tensor, gather_list: Optional[List[torch.Tensor]], dst: int) -> None:
        t, _ = self._contig(tensor)
        flat = torch.empty(self.world * t.numel(), dtype=t.dtype, device=t.device) if self.rank == dst else None
        if self.world == 1:
            gather_list[0].detach().copy_(t)
            return
        with torch.cuda.device(self._torch_dev):
            rc = self.lib.dcp_all_gather(self._h, t.data_ptr(), flat.data_ptr() if flat is not None else None,
                                         t.numel() * t.element_size(), dst, CH_USER, self.default_max_ctas, self._stream())
        self._check(rc, "gather")
        if self.rank == dst:
            for r, o in enumerate(gather_list):
                o.detach().copy_(flat[r * t.numel(): (r + 1) * t.numel()].view(o.shape))

    def scatter(self, tensor: torch.Tensor, scatter_list: Optional[List[torch.Tensor]], src: int) -> None:
        out, back = self._contig(tensor)
        if self.world == 1:
            tensor.detach().copy_(scatter_list[0])
            return
        flat = torch.cat([s.detach().reshape(-1) for s in scatter_list]) if self.rank == src else None
        with torch.cuda.device(self._torch_dev):
            rc = self.lib.dcp_scatter(self._h, flat.data_ptr() if flat is not None else None, out.data_ptr(),
                                      out.numel() * out.element_size(), src, CH_USER, self.default_max_ctas, self._stream())
        self._check(rc, "scatter")
        if back:
            tensor.detach().copy_(out)

    def reduce_scatter_tensor(self, output: torch.Tensor, input: torch.Tensor, op: ReduceOp = ReduceOp.SUM, stream=None) -> None:
        inp, _ = self._contig(input)
        out, back = self._contig(output)
        if inp.numel() != out.numel() * self.world:
            raise ValueError("reduce_scatter_tensor: input must have world_size * output.numel() elements")
        if self.world == 1:
            output.detach().copy_(inp.view(output.shape))
            return
        with torch.cuda.device(self._torch_dev):
            rc = self.lib.dcp_reduce_scatter(self._h, inp.data_ptr(), out.data_ptr(), out.numel(), self._code(out), int(op), -1, CH_USER, self.default_max_ctas,
                                             self._stream(stream))
        self._check(rc, "reduce_scatter")
        if back:
            output.detach().copy_(out)

    def reduce_scatter(self, output: torch.Tensor, input_list: List[torch.Tensor], op: ReduceOp = ReduceOp.SUM) -> None:
        self.reduce_scatter_tensor(output, torch.cat([t.detach().reshape(-1) for t in input_list]), op)

    def send(self, tensor: torch.Tensor, dst: int, tag: int = 0) -> None:
        self._join_async()
        t, _ = self._contig(tensor)
        with torch.cuda.device(self._torch_dev):
            rc = self.lib.dcp_send(self._h, t.data_ptr(), t.numel() * t.element_size(), dst, self._stream())
        self._check(rc, "send")

    def recv(self, tensor: torch.Tensor, src: int, tag: int = 0) -> None:
        self._join_async()
        t, back = self._contig(tensor)
        with torch.cuda.device(self._torch_dev):
            rc = self.lib.dcp_recv(self._h, t.data_ptr(), t.numel() * t.element_size(), src, self._stream())
        self._check(rc, "recv")
        if back:
            tensor.detach().copy_(t)

    # Non-blocking point to point: sends are queued on one side stream and receives on another (each ordered behind the
    # caller's current stream), so a receive kernel that is still waiting for its peer never sits in front of this rank's own
    # send - the exchange "irecv, isend, wait both" that every rank runs at once makes progress.  Operations of one kind stay
    # ordered among themselves; the Work's wait() orders the caller's stream behind the transfer (no host block).
    def _async_stream(self, kind: str):
        st = getattr(self, "_p2p_streams", None)
        if st is None:
            st = self._p2p_streams = {}
        if kind not in st:
            st[kind] = torch.cuda.Stream(self._torch_dev)
        return st[kind]

    def _join_async(self) -> None:
        """Blocking transfers and collectives run behind everything still queued on the two side streams."""
        for s in (getattr(self, "_p2p_streams", None) or {}).values():
            torch.cuda.current_stream(self._torch_dev).wait_stream(s)

    def isend(self, tensor: torch.Tensor, dst: int, tag: int = 0) -> Work:
        t, _ = self._contig(tensor)
        s = self._async_stream("tx")
        s.wait_stream(torch.cuda.current_stream(self._torch_dev))
        with torch.cuda.device(self._torch_dev):
            rc = self.lib.dcp_send(self._h, t.data_ptr(), t.numel() * t.element_size(), dst, s.cuda_stream)
        self._check(rc, "isend")
        t.record_stream(s)
        ev = torch.cuda.Event(); ev.record(s)
        return Work(event=ev, keep=(tensor, t))

    def irecv(self, tensor: torch.Tensor, src: int, tag: int = 0) -> Work:
        t, back = self._contig(tensor)
        s = self._async_stream("rx")
        s.wait_stream(torch.cuda.current_stream(self._torch_dev))
        with torch.cuda.device(self._torch_dev):
            rc = self.lib.dcp_recv(self._h, t.data_ptr(), t.numel() * t.element_size(), src, s.cuda_stream)
        self._check(rc, "irecv")
        if back:
            with torch.cuda.stream(s):
                tensor.detach().copy_(t)
        t.record_stream(s)
        ev = torch.cuda.Event(); ev.record(s)
        return Work(event=ev, keep=(tensor, t))

    def barrier(self) -> None:
        """Device barrier across ranks + host sync (so it also orders host-side effects)."""
        if self.world > 1:
            with torch.cuda.device(self._torch_dev):
                rc = self.lib.dcp_barrier_launch(self._h, CH_USER, self._stream())
            self._check(rc, "barrier")
        torch.cuda.current_stream(self._torch_dev).synchronize()
        self.check_health()

    @property
    def handle(self):
        return self._h

    def close(self) -> None:
        if self._h:
            with torch.cuda.device(self._torch_dev):
                torch.cuda.synchronize()
                self.lib.dcp_symm_destroy(self._h)
            self._h = None
