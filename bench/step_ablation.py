"""What does each piece of the data-parallel machinery cost on the critical path of the flagship step?

Runs the graph-captured fused ConvNet step (parallel/convnet_trainer.py) in several variants inside ONE job and
times each on the device (CUDA events, barrier + synchronize on both sides, max over ranks).  Variants switch one
piece off (DCP_ABLATE) or select the round-1 choreography (DCP_DP_TAIL=0), so the difference to the full step is
that piece's exposed cost.  Ablated variants compute WRONG gradients - they exist only to be timed.

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 \
        bench/step_ablation.py --steps 500 --out profiles/step_ablation_2gpu.json
"""
import argparse, json, os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import distributed_compute_pytorch_b200 as dcp
from distributed_compute_pytorch_b200 import ops
from distributed_compute_pytorch_b200.utils import ClockSampler

VARIANTS = [
    # name, env
    ("full (conv2: local sums + signal, then pull / average / update on a second side stream; conv1: low-latency exchange + update inside its weight-gradient kernel; BN buffers pushed by the head)", {}),
    ("round-1 choreography (side-stream all-reduces, in-kernel conv1 all-reduce, BN broadcast join)", {"DCP_DP_TAIL": "0", "DCP_ABLATE": "oldbn"}),
    ("full, no BN buffer sync", {"DCP_ABLATE": "bn"}),
    ("full, no bucket-0 all-reduce", {"DCP_ABLATE": "b0"}),
    ("full, no bucket-1 exchange", {"DCP_ABLATE": "b1"}),
    ("no communication at all", {"DCP_ABLATE": "bn,b0,b1"}),
    ("full, conv2 reduce + pulling update on the main stream (update after the data-gradient)", {"DCP_ABLATE": "pull_main"}),
    ("full, conv2 sums PUSHED to every rank on the second side stream", {"DCP_ABLATE": "push"}),
    ("full, conv1 through payload + signal + tail kernel instead of the low-latency words", {"DCP_ABLATE": "no_ll"}),
    ("full, 8 comm CTAs", {"DCP_COMM_CTAS": "8"}),
    ("full, 32 comm CTAs", {"DCP_COMM_CTAS": "32"}),
    ("full (repeat)", {}),
]
KEYS = ("DCP_ABLATE", "DCP_DP_TAIL", "DCP_COMM_CTAS", "DCP_DP_NVLS", "DCP_ALLREDUCE_ALGO")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=500); ap.add_argument("--warmup", type=int, default=50)
    ap.add_argument("--batch", type=int, default=128); ap.add_argument("--out", default="")
    opt = ap.parse_args()
    rank, local, world = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1"); os.environ.setdefault("MASTER_PORT", "29611")
    dcp.init_process_group("nccl", rank=rank, world_size=world, device_id=local)
    from distributed_compute_pytorch_b200.parallel.convnet_trainer import FusedConvNetTrainer
    g = torch.Generator().manual_seed(100 + rank)
    pool = 512
    xs = torch.randn(pool, opt.batch, 1, 28, 28, generator=g).to(dev)
    ys = torch.randint(0, 10, (pool, opt.batch), generator=g).to(dev)
    out = []
    for name, env in VARIANTS:
        if world == 1 and env:
            continue
        for k in KEYS:
            os.environ.pop(k, None)
        os.environ.update(env)
        be = dcp.distributed._world().cuda
        if be is not None:      # knobs the native side / backend read once at start-up
            be._algo_override = int(env.get("DCP_ALLREDUCE_ALGO", "0"))
            be.lib.dcp_dp_set_nvls(int(env.get("DCP_DP_NVLS", "0")))
        torch.manual_seed(0)
        model = dcp.models.ConvNet().to(dev).train()
        if world > 1:
            model = dcp.DistributedDataParallel(model, device_ids=[local])
        optim = ops.Adadelta(model.parameters(), lr=1e-3)
        tr = FusedConvNetTrainer(model, optim, opt.batch)

        def run(n):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            dcp.barrier(); torch.cuda.synchronize()
            a.record()
            for i in range(n):
                tr.step(xs[i % pool], ys[i % pool])
            b.record()
            dcp.barrier(); torch.cuda.synchronize()
            return a.elapsed_time(b)
        run(opt.warmup)
        wait_ctr = tr.dp_ws[6:8].view(torch.int64) if getattr(tr, "dp_tail", False) else None     # ns the conv1 exchange spent waiting for peers
        if wait_ctr is not None:
            wait_ctr.zero_()
        with ClockSampler(local, period_s=0.05) as clocks:
            ms = min(run(opt.steps) for _ in range(3))
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        if world > 1:
            dcp.all_reduce(t, op=dcp.ReduceOp.MAX)
        rec = {"variant": name, "env": env, "n_gpus": world, "us_per_step": round(t.item() * 1e3 / opt.steps, 2), "steps": opt.steps,
               "launches_per_step": tr.launches_per_step, "loss": float(tr.loss), "clocks": clocks.summary()}
        if wait_ctr is not None:
            w = wait_ctr.double().clone()
            dcp.all_reduce(w, op=dcp.ReduceOp.MAX)
            rec["tail_wait_us_per_step_max_rank"] = round(w.item() / 1e3 / (3 * opt.steps), 2)
        out.append(rec)
        if rank == 0:
            print(json.dumps(rec), flush=True)
        del tr, optim, model
    if rank == 0 and opt.out:
        json.dump(out, open(opt.out, "w"), indent=1)
    dcp.destroy_process_group()


if __name__ == "__main__":
    main()
