"""North-star model benchmarks (BASELINE.json configs[2..4]): ResNet-18 on CIFAR-shaped data, ResNet-50 bf16 with
SyncBatchNorm on ImageNet-shaped data, GPT-2 small bf16 seq 1024 with the gradient collective fused into the
optimizer step - this framework (`--impl ours`) against the stock PyTorch stack (`--impl torch`: torchvision /
nn.Module models, torch DDP + NCCL, cuDNN / cuBLAS, autocast, torch.optim) on the same box.

    python bench/train_models.py --model resnet18 --gpus 2 --steps 30 --warmup 5 [--impl torch]
    python bench/train_models.py --model resnet50 --bf16 --sync-bn --batch 64 ...
    python bench/train_models.py --model gpt2 --bf16 --batch 8 ...

Synthetic data, random-init weights; device-timed (CUDA events, barrier + synchronize on both sides, max over ranks);
one JSON line on rank 0 with images/s or tokens/s for the whole job.
"""
import argparse
import json
import os
import socket
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

SHAPES = {"resnet18": ((3, 32, 32), 10, 256), "resnet50": ((3, 224, 224), 1000, 64), "gpt2": (None, 50257, 8)}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="resnet18", choices=list(SHAPES))
    ap.add_argument("--impl", default="ours", choices=["ours", "torch"])
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--batch", type=int, default=0, help="per-GPU batch (default: 256 / 64 / 8)")
    ap.add_argument("--seq", type=int, default=1024)
    ap.add_argument("--bf16", action="store_true")
    ap.add_argument("--sync-bn", action="store_true")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--layers", type=int, default=12, help="gpt2 only (12 = GPT-2 small)")
    ap.add_argument("--out", default="", help="append the JSON line to this file as well")
    return ap.parse_args()


def data(opt, dev, rank):
    g = torch.Generator().manual_seed(11 + rank)
    shape, classes, dflt = SHAPES[opt.model]
    B = opt.batch or dflt
    if opt.model == "gpt2":
        tok = torch.randint(0, classes, (4, B, opt.seq + 1), generator=g).to(dev)
        return [(tok[i, :, :-1].contiguous(), tok[i, :, 1:].contiguous()) for i in range(4)], B, B * opt.seq
    xs = torch.randn((4, B) + shape, generator=g).to(dev)
    ys = torch.randint(0, classes, (4, B), generator=g).to(dev)
    return [(xs[i], ys[i]) for i in range(4)], B, B


def run_ours(opt, rank, world):
    import distributed_compute_pytorch_b200 as dcp
    from distributed_compute_pytorch_b200 import ops
    from distributed_compute_pytorch_b200.ops import cuda as K
    from distributed_compute_pytorch_b200.parallel.engine import GraphedTrainStep
    from distributed_compute_pytorch_b200.parallel.zero import FusedShardedOptimizer
    dev = torch.device("cuda", rank)
    torch.cuda.set_device(dev)
    dcp.init_process_group("nccl", rank=rank, world_size=world, device_id=rank)
    torch.manual_seed(0)
    if opt.model == "gpt2":
        model = dcp.models.GPT2(dcp.models.GPT2Config(n_layer=opt.layers)).to(dev)
        loss_fn = None
    else:
        model = (dcp.models.resnet18(10) if opt.model == "resnet18" else dcp.models.resnet50(1000)).to(dev)
        if opt.sync_bn:
            model = ops.SyncBatchNorm.convert_sync_batchnorm(model).to(dev)
        loss_fn = lambda o, t: ops.cross_entropy(o, t)
    if opt.bf16:
        model.compute_dtype = torch.bfloat16
    model.train()
    ddp = dcp.DistributedDataParallel(model, device_ids=[rank], bucket_cap_mb=(256 if opt.model == "gpt2" else None))
    if opt.model == "gpt2":
        optim = FusedShardedOptimizer(ddp, kind="adamw", lr=3e-4, weight_decay=0.1,
                                      shadow_dtype=(torch.bfloat16 if opt.bf16 and os.environ.get("DCP_BF16_SHADOW", "1") == "1" else None))   # reduce_scatter -> AdamW -> all_gather per bucket
    else:
        optim = ops.SGD(ddp.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
    batches, B, units = data(opt, dev, rank)

    if opt.model == "gpt2":
        loss_fn = lambda o, t: ops.cross_entropy(o.view(-1, o.shape[-1]), t.reshape(-1))
    if opt.no_graph:
        def step(x, y):
            optim.zero_grad()
            loss = loss_fn(ddp(x), y)
            loss.backward(); optim.step(); return loss.detach()
    else:
        step = GraphedTrainStep(ddp, optim, loss_fn, batches[0][0], batches[0][1], device=dev, warmup=3)
    return step, batches, units, dcp, (lambda: dcp.barrier()), (lambda: dcp.destroy_process_group()), K


def run_torch(opt, rank, world):
    import torch.distributed as dist
    from torch.nn.parallel import DistributedDataParallel as DDP
    dev = torch.device("cuda", rank)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    torch.manual_seed(0)
    if opt.model == "gpt2":
        os.environ["DCP_FORCE_TORCH_OPS"] = "1"      # same module definition, stock torch ops (cuBLAS, ATen) on CUDA
        import distributed_compute_pytorch_b200 as dcp
        model = dcp.models.GPT2(dcp.models.GPT2Config(n_layer=opt.layers)).to(dev)
    else:
        import torchvision
        model = (torchvision.models.resnet18(num_classes=10) if opt.model == "resnet18" else torchvision.models.resnet50(num_classes=1000)).to(dev)
        model = model.to(memory_format=torch.channels_last)
        if opt.sync_bn:
            model = torch.nn.SyncBatchNorm.convert_sync_batchnorm(model)
    model.train()
    ddp = DDP(model, device_ids=[rank], bucket_cap_mb=(256 if opt.model == "gpt2" else 25))
    if opt.model == "gpt2":
        optim = torch.optim.AdamW(ddp.parameters(), lr=3e-4, weight_decay=0.1, fused=True)
    else:
        optim = torch.optim.SGD(ddp.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
    batches, B, units = data(opt, dev, rank)

    def step(x, y):
        optim.zero_grad(set_to_none=True)
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=opt.bf16):
            if opt.model == "gpt2":
                logits = ddp(x)
                loss = F.cross_entropy(logits.view(-1, logits.shape[-1]).float(), y.reshape(-1))
            else:
                loss = F.cross_entropy(ddp(x.contiguous(memory_format=torch.channels_last)), y)
        loss.backward(); optim.step(); return loss.detach()
    return step, batches, units, dist, (lambda: dist.barrier()), (lambda: dist.destroy_process_group()), None


def worker(rank, world, port, opt):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    step, batches, units, dist, barrier, finish, K = (run_ours if opt.impl == "ours" else run_torch)(opt, rank, world)
    dev = torch.device("cuda", rank)

    def timed(n):
        barrier(); torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for i in range(n):
            loss = step(*batches[i % len(batches)])
        b.record(); barrier(); torch.cuda.synchronize()
        return a.elapsed_time(b), float(loss)
    from distributed_compute_pytorch_b200.utils import ClockSampler
    timed(opt.warmup)
    clocks = ClockSampler(rank, period_s=0.1) if rank == 0 else None     # SM clocks / throttle reasons DURING the timed region
    if clocks:
        clocks.__enter__()
    ms, last = timed(opt.steps)
    if clocks:
        clocks.__exit__(None, None, None)
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = t.item()
    if rank == 0:
        unit = "tokens/s" if opt.model == "gpt2" else "images/s"
        rec = json.dumps({"model": opt.model, "impl": opt.impl, "clocks": clocks.summary(), "warmup": opt.warmup, "n_gpus": world, "value": units * world * opt.steps / (ms / 1e3), "unit": unit,
                          "ms_per_step": ms / opt.steps, "per_gpu_batch": opt.batch or SHAPES[opt.model][2], "bf16": opt.bf16, "sync_bn": opt.sync_bn,
                          "graph": (not opt.no_graph) if opt.impl == "ours" else False, "last_loss": last, "steps": opt.steps,
                          "peak_mem_GB": round(torch.cuda.max_memory_allocated() / 2 ** 30, 2)})
        print(rec, flush=True)
        if opt.out:
            with open(opt.out, "a") as f:
                f.write(rec + "\n")
    finish()


def main():
    opt = parse()
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]
    import distributed_compute_pytorch_b200 as dcp
    dcp.spawn(worker, args=(opt.gpus, port, opt), nprocs=opt.gpus, join=True)


if __name__ == "__main__":
    main()
