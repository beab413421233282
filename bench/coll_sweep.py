"""Collective bandwidth/latency sweep (BASELINE.json configs[1]): ours (NVLink P2P / NVLS kernels) vs
NCCL on the same box, 1 KB - 1 GB, device-timed with CUDA events, max over ranks.

    python bench/coll_sweep.py --gpus 2 [--max-mb 1024] [--ops all_reduce,all_gather,...] [--out gpurun_out/sweep.jsonl]

Bus bandwidth uses the usual conventions: all_reduce 2(W-1)/W * S / t, all_gather / reduce_scatter
(W-1)/W * S / t (S = full buffer), broadcast and send/recv S / t.
"""
import argparse
import json
import os
import socket
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import distributed_compute_pytorch_b200 as dcp  # noqa: E402
from distributed_compute_pytorch_b200.utils import ClockSampler  # noqa: E402

CONVNET_MESSAGES = [4, 8, 1024, 75264, 4725288, 4800552]


def _time(fn, iters, warmup, flush=None):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    start.record()
    for _ in range(iters):
        fn()
    end.record()
    torch.cuda.synchronize()
    return start.elapsed_time(end) / iters * 1e3  # us


def _maxr(x, dev):
    t = torch.tensor([x], device=dev, dtype=torch.float64)
    dcp.all_reduce(t, op=dcp.ReduceOp.MAX)
    return t.item()


def worker(rank, world, port, opt):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    os.environ.setdefault("DCP_HEAP_MB", str(max(2048, 3 * opt.max_mb + 1024)))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dcp.init_process_group("nccl", rank=rank, world_size=world)
    be = dcp.distributed._world().cuda
    import torch.distributed as tdist
    nccl = None
    if not opt.no_nccl:
        tdist.init_process_group("nccl", init_method=f"tcp://127.0.0.1:{port + 1}", rank=rank, world_size=world, device_id=dev)
        nccl = tdist
    sizes = sorted(set([1 << k for k in range(10, 31) if (1 << k) <= opt.max_mb << 20] + CONVNET_MESSAGES))
    sizes = [b for b in sizes if b >= opt.min_kb << 10]
    ops = opt.ops.split(",")
    out = open(opt.out, "a") if rank == 0 and opt.out else None
    dtype = torch.float32
    for nbytes in sizes:
        n = max(nbytes // 4, 1)
        iters = 200 if nbytes <= 1 << 20 else (50 if nbytes <= 64 << 20 else 10)
        warm = 10 if nbytes <= 64 << 20 else 3
        sym = be.empty_symmetric(n, dtype); sym.fill_(1.0)
        loc = torch.ones(n, device=dev, dtype=dtype)
        rows = []
        clocks = ClockSampler(rank, period_s=0.1) if rank == 0 else None     # clocks during this size's timed regions
        if clocks:
            clocks.__enter__()
        if "all_reduce" in ops:
            variants = []
            if nbytes <= be.oneshot_max:
                variants.append(("oneshot", lambda: be.all_reduce(loc, dcp.ReduceOp.SUM, algo=1)))
            variants.append(("twoshot_sym", lambda: be.all_reduce(sym, dcp.ReduceOp.SUM, algo=2)))
            if be.nvls:
                variants.append(("nvls_sym", lambda: be.all_reduce(sym, dcp.ReduceOp.SUM, algo=3)))
            variants.append(("auto", lambda: be.all_reduce(loc, dcp.ReduceOp.SUM)))
            variants.append(("auto_sym", lambda: be.all_reduce(sym, dcp.ReduceOp.SUM)))
            if nccl:
                variants.append(("nccl", lambda: nccl.all_reduce(loc)))
            for name, fn in variants:
                us = _maxr(_time(fn, iters, warm), dev)
                loc.fill_(1.0); sym.fill_(1.0)
                rows.append(("all_reduce", name, us, 2 * (world - 1) / world * nbytes / us / 1e3))
        if "broadcast" in ops:
            variants = [("ours", lambda: be.broadcast(loc, 0)), ("ours_sym", lambda: be.broadcast(sym, 0))]
            if nccl:
                variants.append(("nccl", lambda: nccl.broadcast(loc, 0)))
            for name, fn in variants:
                us = _maxr(_time(fn, iters, warm), dev)
                rows.append(("broadcast", name, us, nbytes / us / 1e3))
        if "all_gather" in ops and n % world == 0 and n >= world:
            shard = torch.ones(n // world, device=dev, dtype=dtype)
            variants = [("ours_sym", lambda: be.all_gather_into_tensor(sym, shard)), ("ours", lambda: be.all_gather_into_tensor(loc, shard))]
            if nccl:
                variants.append(("nccl", lambda: nccl.all_gather_into_tensor(loc, shard)))
            for name, fn in variants:
                us = _maxr(_time(fn, iters, warm), dev)
                rows.append(("all_gather", name, us, (world - 1) / world * nbytes / us / 1e3))
        if "reduce_scatter" in ops and n % world == 0 and n >= world:
            shard = torch.empty(n // world, device=dev, dtype=dtype)
            variants = [("ours_sym", lambda: be.reduce_scatter_tensor(shard, sym)), ("ours", lambda: be.reduce_scatter_tensor(shard, loc))]
            if nccl:
                variants.append(("nccl", lambda: nccl.reduce_scatter_tensor(shard, loc)))
            for name, fn in variants:
                us = _maxr(_time(fn, iters, warm), dev)
                rows.append(("reduce_scatter", name, us, (world - 1) / world * nbytes / us / 1e3))
        if "sendrecv" in ops and world >= 2:
            def ours_sr():
                if rank == 0: be.send(loc, 1)
                elif rank == 1: be.recv(loc, 0)
            def nccl_sr():
                if rank == 0: nccl.send(loc, 1)
                elif rank == 1: nccl.recv(loc, 0)
            variants = [("ours", ours_sr)] + ([("nccl", nccl_sr)] if nccl else [])
            for name, fn in variants:
                us = _maxr(_time(fn, iters, warm), dev)
                rows.append(("sendrecv", name, us, nbytes / us / 1e3))
        be.free_symmetric(sym)
        del loc
        csum = None
        if clocks:
            clocks.__exit__(None, None, None)
            csum = clocks.summary()
        if rank == 0:
            for op, name, us, gbs in rows:
                rec = {"op": op, "impl": name, "bytes": nbytes, "world": world, "us": round(us, 2), "bus_GBps": round(gbs, 2), "nvls": be.nvls,
                       "clocks": csum}
                print(json.dumps(rec), flush=True)
                if out:
                    out.write(json.dumps(rec) + "\n"); out.flush()
    if nccl:
        tdist.destroy_process_group()
    dcp.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=2)
    ap.add_argument("--max-mb", type=int, default=1024)
    ap.add_argument("--min-kb", type=int, default=0, help="skip messages below this size")
    ap.add_argument("--ops", default="all_reduce,broadcast,all_gather,reduce_scatter,sendrecv")
    ap.add_argument("--out", default="")
    ap.add_argument("--no-nccl", action="store_true")
    opt = ap.parse_args()
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]
    if opt.out:
        os.makedirs(os.path.dirname(opt.out) or ".", exist_ok=True)
    dcp.spawn(worker, args=(opt.gpus, port, opt), nprocs=opt.gpus, join=True)


if __name__ == "__main__":
    main()
