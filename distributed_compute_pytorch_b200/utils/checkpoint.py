"""Checkpoint helpers.  The reference saves once, model only: ``torch.save(model.state_dict(), "mnist.pt")``
(ref:main.py:133) from EVERY rank at once.  Here:

* ``save_model(path, model)`` writes exactly that file (an ``OrderedDict`` of per-parameter tensors, ``module.*`` keys when the
  model is DDP-wrapped) - from rank 0 only, through a temporary name + rename, followed by a barrier: no concurrent writers.
* ``save_training_state`` / ``load_training_state`` add what the reference lacks for resuming (optimizer, LR schedule, epoch) in
  a SEPARATE file, so ``mnist.pt`` stays loadable by the reference's ``ConvNet`` / by plain ``load_state_dict``.
Parameters that live as views into flat buckets come out as ordinary per-parameter tensors (``state_dict()`` clones to CPU).
"""
from __future__ import annotations

import os
from typing import Optional

import torch


def _rank0() -> bool:
    from .. import distributed as dist
    return (not dist.is_initialized()) or dist.get_rank() == 0


def _barrier() -> None:
    from .. import distributed as dist
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.barrier()


def _atomic_save(obj, path: str) -> None:
    tmp = f"{path}.tmp.{os.getpid()}"
    torch.save(obj, tmp)
    os.replace(tmp, path)


def save_model(path: str, model: torch.nn.Module) -> None:
    """The reference's checkpoint: ``model.state_dict()`` on CPU, written by rank 0, every rank waits until it exists."""
    if _rank0():
        sd = model.state_dict()
        out = type(sd)((k, v.detach().cpu().clone()) for k, v in sd.items())
        if hasattr(sd, "_metadata"):
            out._metadata = sd._metadata
        _atomic_save(out, path)
    _barrier()


def save_training_state(path: str, optimizer=None, scheduler=None, epoch: Optional[int] = None, extra: Optional[dict] = None) -> None:
    if _rank0():
        state = {"epoch": epoch, "extra": extra or {}}
        if optimizer is not None:
            state["optimizer"] = optimizer.state_dict()
        if scheduler is not None:
            state["scheduler"] = scheduler.state_dict()
        _atomic_save(state, path)
    _barrier()


def load_model(path: str, model: torch.nn.Module, strict: bool = True) -> None:
    """Load a reference-format checkpoint; ``module.`` prefixes are added / stripped to match ``model`` (DDP-wrapped or not)."""
    sd = torch.load(path, map_location="cpu")
    want_prefix = any(k.startswith("module.") for k in model.state_dict().keys())
    has_prefix = any(k.startswith("module.") for k in sd.keys())
    if want_prefix and not has_prefix:
        sd = type(sd)(("module." + k, v) for k, v in sd.items())
    elif has_prefix and not want_prefix:
        sd = type(sd)((k[len("module."):] if k.startswith("module.") else k, v) for k, v in sd.items())
    model.load_state_dict(sd, strict=strict)


def load_training_state(path: str, optimizer=None, scheduler=None) -> dict:
    state = torch.load(path, map_location="cpu")
    if optimizer is not None and "optimizer" in state:
        optimizer.load_state_dict(state["optimizer"])
    if scheduler is not None and "scheduler" in state:
        scheduler.load_state_dict(state["scheduler"])
    return state
