"""Headline benchmark: images/sec of the reference model's DDP training step (BASELINE.json metric;
BASELINE.md row 2: ConvNet, synthetic MNIST shape [B,1,28,28], batch 128/GPU, fp32, weak scaling).

    python bench.py --gpus N --steps K --warmup W                  # this framework
    python bench.py --impl reference --gpus N --steps K --warmup W # unmodified reference (baseline/_ref) on torch DDP + NCCL

For N > 1 launch under torchrun (one rank per GPU); without torchrun env the script spawns its own ranks.
One JSON line on rank 0.  `value` = device-timed steps on device-resident inputs (CUDA events, barrier +
synchronize on both sides, max over ranks); `e2e` = the same metric through the public API including, every
step, the H2D copy of the step's inputs from pinned host memory and a D2H read of the loss.
"""
import argparse
import contextlib
import importlib.util
import io
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

POOL_BATCHES = 512  # rotating input pool: 512 x 128 x 784 x 4 B = 205 MB > 126 MB L2


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=1000)
    ap.add_argument("--warmup", type=int, default=50)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "reference-graphed"],
                    help="reference-graphed is a DIAGNOSTIC arm (not the headline comparison): the reference's ConvNet and the body of its "
                         "train() loop captured with torch.cuda.graphs, to separate kernel quality from eager launch overhead")
    ap.add_argument("--batch", type=int, default=128, help="per-GPU batch (the reference's --batch_size default, ref:main.py:139)")
    ap.add_argument("--lr", type=float, default=0.001)
    ap.add_argument("--no-graph", action="store_true", help="ours: eager DDP path instead of the CUDA-graph step")
    ap.add_argument("--engine", default="fused", choices=["fused", "autograd"],
                    help="ours: fused = parallel.FusedConvNetTrainer (flat-shift tcgen05 convs); autograd = layer-by-layer ops under GraphedTrainStep")
    return ap.parse_args()


def env_rank():
    if "RANK" in os.environ and "WORLD_SIZE" in os.environ:
        return int(os.environ["RANK"]), int(os.environ.get("LOCAL_RANK", os.environ["RANK"])), int(os.environ["WORLD_SIZE"])
    return None


def make_pool(batch, device, pinned, seed):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(POOL_BATCHES, batch, 1, 28, 28, generator=g)
    y = torch.randint(0, 10, (POOL_BATCHES, batch), generator=g)
    if pinned:
        return x.pin_memory(), y.pin_memory()
    return x.to(device), y.to(device)


def max_over_ranks(val, device, dist_mod, world):
    if world == 1:
        return val
    t = torch.tensor([val], dtype=torch.float64, device=device)
    dist_mod.all_reduce(t, op=dist_mod.ReduceOp.MAX)
    return t.item()


# ----------------------------------------------------------------------------------------- ours ----
def run_ours(opt, rank, local_rank, world):
    import distributed_compute_pytorch_b200 as dcp
    from distributed_compute_pytorch_b200 import ops
    from distributed_compute_pytorch_b200.models import ConvNet
    from distributed_compute_pytorch_b200.ops import cuda as K
    from distributed_compute_pytorch_b200.parallel.engine import GraphedTrainStep
    from distributed_compute_pytorch_b200.utils import ClockSampler

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29533")
    dcp.init_process_group("nccl", rank=rank, world_size=world, device_id=local_rank)
    be = dcp.distributed._world().cuda
    torch.manual_seed(0)
    model = ConvNet().to(dev)
    if world > 1:
        model = dcp.DistributedDataParallel(model, device_ids=[local_rank])
    optimizer = ops.Adadelta(model.parameters(), lr=opt.lr)
    model.train()
    xs, ys = make_pool(opt.batch, dev, pinned=False, seed=100 + rank)
    hx, hy = make_pool(opt.batch, dev, pinned=True, seed=100 + rank)

    if opt.engine == "fused":
        from distributed_compute_pytorch_b200.parallel.convnet_trainer import FusedConvNetTrainer
        step = FusedConvNetTrainer(model, optimizer, opt.batch, use_graph=not opt.no_graph)
        per_step = step.launches_per_step if not opt.no_graph else None
    elif opt.no_graph:
        def step(x, y):
            optimizer.zero_grad()
            loss = ops.nll_loss(model(x.to(dev, non_blocking=True)), y.to(dev, non_blocking=True))
            loss.backward()
            optimizer.step()
            return loss.detach()
        per_step = None
    else:
        step = GraphedTrainStep(model, optimizer, lambda o, t: ops.nll_loss(o, t), xs[0], ys[0], device=dev)
        per_step = step.launches_per_step

    def timed(pool_x, pool_y, nsteps, read_loss):
        # end-to-end mode: every step's loss is copied to pinned host memory and READ on the host inside the timed
        # region; the read of step i-1 happens after step i has been enqueued (two slots + events), the way a training
        # loop logs without draining the GPU every step
        loss_host = torch.zeros(2, dtype=torch.float32).pin_memory()
        done = [torch.cuda.Event(), torch.cuda.Event()]
        start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        dcp.barrier(); torch.cuda.synchronize()
        l0 = K.launch_count + be.launches
        t0 = time.perf_counter()
        start.record()
        last = 0.0
        for i in range(nsteps):
            loss = step(pool_x[i % POOL_BATCHES], pool_y[i % POOL_BATCHES])
            if read_loss:
                loss_host[i & 1].copy_(loss, non_blocking=True)
                done[i & 1].record()
                if i > 0:
                    done[(i - 1) & 1].synchronize()
                    last = float(loss_host[(i - 1) & 1])
        if read_loss and nsteps > 0:
            done[(nsteps - 1) & 1].synchronize()
            last = float(loss_host[(nsteps - 1) & 1])
        end.record()
        dcp.barrier(); torch.cuda.synchronize()
        wall = time.perf_counter() - t0
        eager_launches = K.launch_count + be.launches - l0
        return start.elapsed_time(end), wall, last, eager_launches

    timed(xs, ys, opt.warmup, False)
    with ClockSampler(local_rank, period_s=0.05) as clocks:
        ms, wall, _, eager_l = timed(xs, ys, opt.steps, False)
    timed(hx, hy, max(3, opt.warmup // 2), True)
    ms_e2e, wall_e2e, last_loss, _ = timed(hx, hy, opt.steps, True)
    ms = max_over_ranks(ms, dev, dcp, world)
    ms_e2e = max_over_ranks(ms_e2e, dev, dcp, world)
    gb = opt.batch * world
    launches = per_step * opt.steps if per_step is not None else eager_l
    rec = {
        "metric": "images/sec", "value": gb * opt.steps / (ms / 1e3), "unit": "images/s", "n_gpus": world, "steps": opt.steps,
        "warmup": opt.warmup, "ms_per_step": ms / opt.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "fp32 (tf32 tensor-core multiply where tcgen05 kernels are used, the precision cuDNN's default gives the reference)",
        "data": "synthetic MNIST-shaped [B,1,28,28] float32 + int64 labels, random-init weights", "impl": "ours",
        "config": {"model": "ConvNet (ref:main.py:20-45)", "global_batch": gb, "per_gpu_batch": opt.batch, "seq_len": None,
                   "parallelism": f"dp{world}", "optimizer": "Adadelta lr=0.001", "graph": not opt.no_graph, "engine": opt.engine,
                   "l2": f"inputs larger than L2: rotating pool of {POOL_BATCHES} batches ({POOL_BATCHES * opt.batch * 784 * 4 / 1e6:.0f} MB)",
                   "nvls": be.nvls},
        "clocks": clocks.summary(),
        "e2e": {"value": gb * opt.steps / (ms_e2e / 1e3), "unit": "images/s", "ms_per_step": ms_e2e / opt.steps,
                "h2d_bytes_per_step": opt.batch * 784 * 4 + opt.batch * 8, "d2h_bytes_per_step": 4, "wall_ms_per_step": wall_e2e * 1e3 / opt.steps,
                "last_loss": last_loss},
        "gpu_launches": launches, "wall_ms_per_step": wall * 1e3 / opt.steps,
    }
    if rank == 0:
        print(json.dumps(rec), flush=True)
    dcp.destroy_process_group()


# ------------------------------------------------------------------------------------ reference ----
def load_reference():
    ref_main = os.path.join(ROOT, "baseline", "_ref", "main.py")
    if not os.path.exists(ref_main):
        sys.path.insert(0, os.path.join(ROOT, "baseline"))
        try:
            import install_ref
            with contextlib.redirect_stdout(io.StringIO()):
                install_ref.install()
        except Exception:
            pass
    if not os.path.exists(ref_main):
        return None, "baseline/_ref/main.py missing and /root/reference not reachable from this box"
    try:
        spec = importlib.util.spec_from_file_location("reference_main", ref_main)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        return mod, None
    except Exception as e:  # e.g. torchvision missing
        return None, f"importing the reference failed: {type(e).__name__}: {e}"


class _Loader:
    """Stands in for the MNIST DataLoader (no network): yields `n` (img, label) batches from a pool."""

    def __init__(self, xs, ys, n):
        self.xs, self.ys, self.n = xs, ys, n
        self.dataset = range(n * xs.shape[1])

    def __len__(self):
        return self.n

    def __iter__(self):
        for i in range(self.n):
            yield self.xs[i % POOL_BATCHES], self.ys[i % POOL_BATCHES]


def run_reference(opt, rank, local_rank, world):
    ref, why = load_reference()
    if ref is None:
        if rank == 0:
            print(json.dumps({"impl": "reference", "unavailable": why}), flush=True)
        return
    import torch.distributed as tdist
    from torch import optim
    from distributed_compute_pytorch_b200.utils import ClockSampler
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29534")
    # deviation 1: the reference hard-codes "gloo" (ref:main.py:50); the bar is its NCCL build
    tdist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    torch.manual_seed(0)                                   # ref:main.py:103
    model = ref.ConvNet().to(local_rank)                   # ref:main.py:118
    if torch.cuda.device_count() > 1:                      # ref:main.py:119
        model = ref.DistributedDataParallel(model, device_ids=[local_rank])   # ref:main.py:122
    optimizer = optim.Adadelta(model.parameters(), lr=opt.lr)                 # ref:main.py:124
    xs, ys = make_pool(opt.batch, dev, pinned=False, seed=100 + rank)
    hx, hy = make_pool(opt.batch, dev, pinned=True, seed=100 + rank)
    ns = argparse.Namespace(batch_size=opt.batch, lr=opt.lr)

    def timed(pool_x, pool_y, nsteps):
        loader = _Loader(pool_x, pool_y, nsteps)
        start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        tdist.barrier(); torch.cuda.synchronize()
        t0 = time.perf_counter()
        start.record()
        with contextlib.redirect_stdout(io.StringIO()):    # the reference prints a loss line every 10 batches
            ref.train(ns, model, local_rank, loader, optimizer, 0)      # the reference's own hot loop, ref:main.py:55-68
        end.record()
        tdist.barrier(); torch.cuda.synchronize()
        return start.elapsed_time(end), time.perf_counter() - t0

    def maxr(v):
        t = torch.tensor([v], dtype=torch.float64, device=dev)
        tdist.all_reduce(t, op=tdist.ReduceOp.MAX)
        return t.item()

    timed(xs, ys, opt.warmup)
    with ClockSampler(local_rank, period_s=0.05) as clocks:
        ms, wall = timed(xs, ys, opt.steps)
    timed(hx, hy, max(3, opt.warmup // 2))
    ms_e2e, wall_e2e = timed(hx, hy, opt.steps)
    ms, ms_e2e = maxr(ms), maxr(ms_e2e)
    gb = opt.batch * world
    rec = {
        "metric": "images/sec", "value": gb * opt.steps / (ms / 1e3), "unit": "images/s", "n_gpus": world, "steps": opt.steps,
        "warmup": opt.warmup, "ms_per_step": ms / opt.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "fp32 (cuDNN default allow_tf32 for conv)", "data": "synthetic MNIST-shaped [B,1,28,28] float32 + int64 labels, random-init weights",
        "impl": "reference",
        "config": {"model": "ConvNet (baseline/_ref/main.py, unmodified)", "global_batch": gb, "per_gpu_batch": opt.batch, "seq_len": None,
                   "parallelism": f"dp{world}", "optimizer": "torch.optim.Adadelta lr=0.001", "stack": "reference train() + torch DDP + NCCL + cuDNN/cuBLAS",
                   "deviations": "backend gloo->nccl; MNIST files -> synthetic tensors",
                   "l2": f"inputs larger than L2: rotating pool of {POOL_BATCHES} batches"},
        "clocks": clocks.summary(),
        "e2e": {"value": gb * opt.steps / (ms_e2e / 1e3), "unit": "images/s", "ms_per_step": ms_e2e / opt.steps,
                "h2d_bytes_per_step": opt.batch * 784 * 4 + opt.batch * 8, "d2h_bytes_per_step": 0.4,
                "note": "reference reads the loss on rank 0 every 10th batch (ref:main.py:64-68)", "wall_ms_per_step": wall_e2e * 1e3 / opt.steps},
        "gpu_launches": 0, "wall_ms_per_step": wall * 1e3 / opt.steps,
    }
    if rank == 0:
        print(json.dumps(rec), flush=True)
    tdist.destroy_process_group()


def run_reference_graphed(opt, rank, local_rank, world):
    """Diagnostic: the reference model + the body of its training loop (ref:main.py:58-63) replayed as ONE CUDA graph on the stock
    PyTorch stack (cuDNN / cuBLAS / ATen, torch.optim.Adadelta capturable, torch DDP + NCCL when world > 1).  The reference itself
    runs eagerly; this arm only answers "how much of the gap is launch overhead"."""
    ref, why = load_reference()
    if ref is None:
        if rank == 0:
            print(json.dumps({"impl": "reference-graphed", "unavailable": why}), flush=True)
        return
    import torch.distributed as tdist
    import torch.nn.functional as F
    from torch import optim
    from distributed_compute_pytorch_b200.utils import ClockSampler
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29535")
    tdist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    torch.manual_seed(0)
    model = ref.ConvNet().to(local_rank)
    if world > 1:
        model = ref.DistributedDataParallel(model, device_ids=[local_rank])
    optimizer = optim.Adadelta(model.parameters(), lr=opt.lr, capturable=True)
    xs, ys = make_pool(opt.batch, dev, pinned=False, seed=100 + rank)
    sx, sy = xs[0].clone(), ys[0].clone()

    def body():
        optimizer.zero_grad(set_to_none=True)
        loss = F.nll_loss(model(sx), sy)
        loss.backward()
        optimizer.step()
        return loss
    try:
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side), warnings_off():
            for _ in range(12):          # DDP needs >= 11 eager iterations before capture
                body()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph), warnings_off():
            static_loss = body()
    except Exception as e:
        if rank == 0:
            print(json.dumps({"impl": "reference-graphed", "unavailable": f"capture failed: {type(e).__name__}: {str(e)[:200]}"}), flush=True)
        tdist.destroy_process_group()
        return

    def timed(n):
        start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        tdist.barrier(); torch.cuda.synchronize()
        start.record()
        for i in range(n):
            sx.copy_(xs[i % POOL_BATCHES], non_blocking=True); sy.copy_(ys[i % POOL_BATCHES], non_blocking=True)
            graph.replay()
        end.record()
        tdist.barrier(); torch.cuda.synchronize()
        return start.elapsed_time(end)
    timed(opt.warmup)
    with ClockSampler(local_rank, period_s=0.05) as clocks:
        ms = timed(opt.steps)
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    tdist.all_reduce(t, op=tdist.ReduceOp.MAX)
    ms = t.item()
    gb = opt.batch * world
    if rank == 0:
        print(json.dumps({"metric": "images/sec", "value": gb * opt.steps / (ms / 1e3), "unit": "images/s", "n_gpus": world, "steps": opt.steps,
                          "warmup": opt.warmup, "ms_per_step": ms / opt.steps, "higher_is_better": True, "scaling": "weak", "impl": "reference-graphed",
                          "diagnostic": True, "dtype": "fp32 (cuDNN default allow_tf32 for conv)", "data": "synthetic MNIST-shaped, random-init weights",
                          "config": {"model": "ConvNet (baseline/_ref/main.py) + body of its train() loop in one torch.cuda.graphs capture",
                                     "global_batch": gb, "per_gpu_batch": opt.batch, "parallelism": f"dp{world}",
                                     "optimizer": "torch.optim.Adadelta(capturable=True)", "l2": f"rotating pool of {POOL_BATCHES} batches"},
                          "clocks": clocks.summary(), "last_loss": float(static_loss)}), flush=True)
    tdist.destroy_process_group()


@contextlib.contextmanager
def warnings_off():
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        yield


def _pick(opt):
    return {"reference": run_reference, "reference-graphed": run_reference_graphed}.get(opt.impl, run_ours)


def _entry(rank, world, port, opt):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    os.environ["RANK"], os.environ["LOCAL_RANK"], os.environ["WORLD_SIZE"] = str(rank), str(rank), str(world)
    _pick(opt)(opt, rank, rank, world)


def main():
    opt = parse()
    er = env_rank()
    if er is not None:
        rank, local_rank, world = er
        _pick(opt)(opt, rank, local_rank, world)
        return
    if opt.gpus <= 1:
        os.environ.setdefault("MASTER_PORT", str(29500 + os.getpid() % 400))
        _pick(opt)(opt, 0, 0, 1)
        return
    import socket
    import distributed_compute_pytorch_b200 as dcp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]
    dcp.spawn(_entry, args=(opt.gpus, port, opt), nprocs=opt.gpus, join=True)


if __name__ == "__main__":
    main()
