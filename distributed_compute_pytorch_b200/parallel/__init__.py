"""Parallelism: DDP wrapper + bucket reducer, NVLink P2P backend, sharded fused optimizer step."""
from .ddp import DistributedDataParallel, broadcast_coalesced, plan_buckets  # noqa: F401
