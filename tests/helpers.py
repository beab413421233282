"""Shared helpers for multi-process tests (module-level functions so 'spawn' can pickle them)."""
import os
import socket


def free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def init(rank: int, world: int, port: int, backend=None):
    import distributed_compute_pytorch_b200 as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group(backend, rank=rank, world_size=world, timeout=120)
    return dist
