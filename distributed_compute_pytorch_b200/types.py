"""Shared enums / handles of the torch.distributed-shaped API."""
from __future__ import annotations

import enum
from typing import Callable, Optional


class ReduceOp(enum.IntEnum):
    """Reduction operators.  The reference only ever uses ``ReduceOp.SUM`` (ref:main.py:65,90,91)."""
    SUM = 0
    PRODUCT = 1
    MIN = 2
    MAX = 3
    BAND = 4
    BOR = 5
    BXOR = 6
    AVG = 7


class Work:
    """Handle returned by ``async_op=True`` collectives and by isend/irecv.

    CUDA collectives are stream-ordered: ``wait()`` makes the caller's current stream wait for the
    collective's stream (no host block), matching the behaviour users expect from NCCL work objects.
    """

    def __init__(self, event=None, stream=None, finalize: Optional[Callable[[], None]] = None, done: bool = True,
                 poll: Optional[Callable[[], bool]] = None, keep=None):
        self._event, self._stream, self._finalize, self._done = event, stream, finalize, done
        self._poll = poll            # host-side operations in flight (CPU isend / irecv): True once finished
        self._keep = keep            # tensors that must outlive the operation

    def wait(self, timeout=None) -> bool:
        if self._finalize is not None:
            self._finalize()
            self._finalize = None
        if self._event is not None:
            import torch
            torch.cuda.current_stream().wait_event(self._event)
        self._done = True
        return True

    def is_completed(self) -> bool:
        if self._poll is not None and self._finalize is not None:
            return bool(self._poll())
        if self._event is not None:
            return bool(self._event.query())
        return self._done and self._finalize is None

    def is_success(self) -> bool:
        return self.is_completed()

    def synchronize(self) -> None:
        self.wait()
        if self._event is not None:
            self._event.synchronize()
