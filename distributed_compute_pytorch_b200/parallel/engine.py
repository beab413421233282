"""CUDA-graph training step: the whole forward + loss + backward + bucket all-reduce + optimizer update
of one iteration of the reference's hot loop (ref:main.py:57-63) captured once and replayed as a single
graph launch - streams and graphs instead of a tracing compiler.

The reference's step is launch-bound on a B200 (about 9 GFLOP of math per 128 images but >100 kernel
launches through the eager stack); replaying one graph removes the host from the loop.  Everything
inside is graph-safe by construction: our kernels take their hyper-parameters (lr, step, dropout
stream) from device memory, the P2P collectives use monotonically increasing epochs, and the DDP
side-stream fork/join is captured as graph dependencies.
"""
from __future__ import annotations

from typing import Callable, Optional

import torch

from ..ops import cuda as K


class GraphedTrainStep:
    """``step = GraphedTrainStep(model, optimizer, loss_fn, x, y); loss = step(x, y)``.

    ``model`` may be a plain module or a ``DistributedDataParallel`` wrapper; ``loss_fn(out, target)``
    must return a 0-dim tensor.  Inputs are copied into static buffers (that copy is the H2D transfer when
    the caller hands pinned host tensors); the returned loss is a static device tensor.
    """

    def __init__(self, model, optimizer, loss_fn: Callable, example_x: torch.Tensor, example_y: torch.Tensor,
                 device: Optional[torch.device] = None, warmup: int = 3):
        self.model, self.optimizer, self.loss_fn = model, optimizer, loss_fn
        dev = device or next(model.parameters()).device
        self.device = dev
        self.x = torch.empty(example_x.shape, dtype=example_x.dtype, device=dev)
        self.y = torch.empty(example_y.shape, dtype=example_y.dtype, device=dev)
        self.x.copy_(example_x); self.y.copy_(example_y)
        if K.graph_step is None:
            K.graph_step = torch.zeros((), dtype=torch.int64, device=dev)
        self.loss = None
        self.launches_per_step = 0
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(warmup):  # bucket rebuild, optimizer state, allocator warm-up happen here
                self._one_step()
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        for p in model.parameters():
            p.grad = None
        from .. import distributed as dist
        be = dist._world().cuda if dist.is_initialized() else None
        before = K.launch_count + (be.launches if be else 0)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.loss = self._one_step()
        self.launches_per_step = K.launch_count + (be.launches if be else 0) - before
        torch.cuda.synchronize(dev)

    def _one_step(self) -> torch.Tensor:
        K.advance_graph_step()
        self.optimizer.zero_grad(set_to_none=True)
        out = self.model(self.x)
        loss = self.loss_fn(out, self.y)
        loss.backward()
        self.optimizer.step()
        return loss.detach()

    def __call__(self, x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
        self.x.copy_(x, non_blocking=True)
        self.y.copy_(y, non_blocking=True)
        if hasattr(self.optimizer, "_hyper"):
            K.sync_optimizer_hyper(self.optimizer)     # StepLR between replays: lr lives in device memory
        self.graph.replay()
        return self.loss
