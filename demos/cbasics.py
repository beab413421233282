"""cbasics: every collective primitive of the framework in one small demo (BASELINE.json configs[1]).

    python demos/cbasics.py --gpus 2            # NVLink P2P backend on CUDA tensors
    python demos/cbasics.py --gpus 2 --cpu      # CPU socket mesh (what "gloo" is to the reference, ref:main.py:50)

Each rank prints what it received; rank 0 ends with "cbasics: all primitives OK".  The bandwidth sweep against NCCL
lives in bench/coll_sweep.py.
"""
import argparse
import os
import socket
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import distributed_compute_pytorch_b200 as dist  # noqa: E402


def run(rank, world, port, cpu):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    if not cpu:
        torch.cuda.set_device(rank)
    dist.init_process_group("gloo" if cpu else "nccl", rank=rank, world_size=world)
    dev = torch.device("cpu") if cpu else torch.device("cuda", rank)
    say = lambda *a: print(f"[rank {rank}]", *a, flush=True)

    # point to point: ring
    nxt, prv = (rank + 1) % world, (rank - 1) % world
    buf = torch.empty(4, device=dev)
    if rank % 2 == 0:
        dist.send(torch.full((4,), float(rank), device=dev), dst=nxt); dist.recv(buf, src=prv)
    else:
        dist.recv(buf, src=prv); dist.send(torch.full((4,), float(rank), device=dev), dst=nxt)
    say("send/recv      got", buf.tolist(), "from rank", prv)
    assert buf[0].item() == prv

    t = torch.arange(4, dtype=torch.float32, device=dev) * (rank + 1)
    dist.broadcast(t, src=0); say("broadcast      ", t.tolist()); assert t.tolist() == [0, 1, 2, 3]

    t = torch.full((4,), float(rank + 1), device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.SUM); say("all_reduce SUM ", t.tolist()); assert t[0].item() == world * (world + 1) / 2

    t = torch.full((4,), float(rank + 1), device=dev)
    dist.reduce(t, dst=0, op=dist.ReduceOp.SUM)
    if rank == 0:
        say("reduce -> 0    ", t.tolist()); assert t[0].item() == world * (world + 1) / 2

    outs = [torch.empty(2, device=dev) for _ in range(world)]
    dist.all_gather(outs, torch.full((2,), float(rank), device=dev)); say("all_gather     ", [o[0].item() for o in outs])
    assert [o[0].item() for o in outs] == list(range(world))

    gl = [torch.empty(2, device=dev) for _ in range(world)] if rank == 0 else None
    dist.gather(torch.full((2,), float(rank * 10), device=dev), gl, dst=0)
    if rank == 0:
        say("gather -> 0    ", [g[0].item() for g in gl]); assert [g[0].item() for g in gl] == [10.0 * r for r in range(world)]

    sl = [torch.full((2,), float(100 + r), device=dev) for r in range(world)] if rank == 0 else None
    t = torch.empty(2, device=dev); dist.scatter(t, sl, src=0); say("scatter <- 0   ", t.tolist()); assert t[0].item() == 100 + rank

    out = torch.empty(2, device=dev)
    dist.reduce_scatter(out, [torch.full((2,), float(rank + r), device=dev) for r in range(world)])
    say("reduce_scatter ", out.tolist()); assert out[0].item() == sum(r + rank for r in range(world))

    dist.barrier()
    if rank == 0:
        print("cbasics: all primitives OK", flush=True)
    dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=2)
    ap.add_argument("--cpu", action="store_true")
    opt = ap.parse_args()
    cpu = opt.cpu or not torch.cuda.is_available()
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]
    dist.spawn(run, args=(opt.gpus, port, cpu), nprocs=opt.gpus, join=True)


if __name__ == "__main__":
    main()
