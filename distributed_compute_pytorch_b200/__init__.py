"""distributed_compute_pytorch_b200 - B200-native data-parallel training + collectives.

A from-scratch framework with the capabilities of ``saandeepa93/distributed_compute_pytorch``
(a PyTorch DDP MNIST example, ``/root/reference/main.py``): a ``torch.distributed``-shaped API,
a ``DistributedDataParallel`` wrapper with ``DistributedSampler`` sharding, and the model/ops/optimizer
stack underneath, re-designed for Blackwell: hand-written sm_100a kernels for the collectives (NVLink
peer memory, no NCCL) and for the compute hot path (tcgen05/TMEM/TMA), CUDA streams + graphs for
scheduling.

Typical use mirrors the reference (ref:main.py:47-53,98-134)::

    import distributed_compute_pytorch_b200 as dcp
    dcp.init_process_group("nccl", rank=rank, world_size=world_size)
    model = dcp.DistributedDataParallel(dcp.models.ConvNet().to(rank), device_ids=[rank])
    sampler = dcp.DistributedSampler(dataset, num_replicas=world_size, rank=rank)
    ...
    dcp.all_reduce(loss, op=dcp.ReduceOp.SUM)
    dcp.destroy_process_group()
"""
from .types import ReduceOp, Work  # noqa: F401
from .distributed import (  # noqa: F401
    init_process_group, destroy_process_group, is_initialized, is_available, get_rank, get_world_size, get_backend,
    send, recv, isend, irecv, broadcast, all_reduce, reduce, all_gather, all_gather_into_tensor, gather, scatter,
    reduce_scatter, reduce_scatter_tensor, barrier, broadcast_object_list, group,
)
from .store import TCPStore, PrefixStore  # noqa: F401
from .launch import spawn, launch, start_processes, ProcessRaisedException, ProcessExitedException  # noqa: F401
from .sampler import DistributedSampler  # noqa: F401
from .parallel.ddp import DistributedDataParallel  # noqa: F401
from . import models, ops, parallel, utils  # noqa: F401

__version__ = "0.1.0"
