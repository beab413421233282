"""Timeline of ONE fused ConvNet training step, on 1 or N GPUs: every launch of FusedConvNetTrainer._issue() is followed by a
CUDA event on the stream it went to (main stream or the communication side stream), the step is queued behind a long memset
so the host never starves the GPU, and each event's offset from the start of the step gives when that kernel finished
under the conditions of the real step - warm L2, predecessor just finished, side stream running alongside.  The third of
three queued steps is reported, from rank 0 and from the slowest rank.  (The event records cost a little and the launches are
not graph-replayed; bench.py's step time is the authoritative total.  This tool shows WHERE the time goes.)

    python bench/step_timeline.py [--batch 128] [--out profiles/step_timeline.json]
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29612 bench/step_timeline.py --out ...
"""
import argparse, json, linecache, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import distributed_compute_pytorch_b200 as dcp
from distributed_compute_pytorch_b200 import ops
from distributed_compute_pytorch_b200.ops import cuda as K
from distributed_compute_pytorch_b200.parallel.convnet_trainer import FusedConvNetTrainer


def main():
    ap = argparse.ArgumentParser(); ap.add_argument("--batch", type=int, default=128); ap.add_argument("--out", default="")
    opt = ap.parse_args()
    rank, local, world = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1"); os.environ.setdefault("MASTER_PORT", "29612")
        dcp.init_process_group("nccl", rank=rank, world_size=world, device_id=local)
    torch.manual_seed(0)
    net = dcp.models.ConvNet().to(dev).train()
    model = dcp.DistributedDataParallel(net, device_ids=[local]) if world > 1 else net
    optim = ops.Adadelta(model.parameters(), lr=1e-3)
    tr = FusedConvNetTrainer(model, optim, opt.batch, use_graph=False)
    tr.x.normal_(); tr.y.random_(0, 10)
    for _ in range(3): tr._issue()
    torch.cuda.synchronize()
    marks = []
    main_stream = torch.cuda.current_stream(dev)
    orig = K._chk

    def src_line():
        g = sys._getframe(2)
        while g is not None and not g.f_code.co_filename.endswith("convnet_trainer.py"):
            g = g.f_back
        if g is None:
            return "?"
        return linecache.getline(g.f_code.co_filename, g.f_lineno).strip()

    def chk(rc):
        orig(rc)
        src = src_line()
        on_side = "opt_stream.cuda_stream" in src and tr.side is not main_stream and "opt0main" not in tr.ablate
        st = tr.side if on_side else main_stream
        e = torch.cuda.Event(enable_timing=True); e.record(st)
        marks.append((src[:100], "side" if on_side else "main", e))
    K._chk = chk
    if tr.be is not None:
        be = tr.be
        orig_ar, orig_bc = be.all_reduce, be.broadcast

        def all_reduce(t, op=None, stream=None, **kw):
            orig_ar(t, op, stream=stream, **kw)
            st = stream or main_stream
            e = torch.cuda.Event(enable_timing=True); e.record(st)
            marks.append((f"all_reduce {t.numel() * t.element_size()} B", "side" if st is not main_stream else "main", e))

        def broadcast(t, src, stream=None, **kw):
            orig_bc(t, src, stream=stream, **kw)
            st = stream or main_stream
            e = torch.cuda.Event(enable_timing=True); e.record(st)
            marks.append((f"broadcast {t.numel() * t.element_size()} B", "side" if st is not main_stream else "main", e))
        be.all_reduce, be.broadcast = all_reduce, broadcast
    flush = torch.empty(1 << 30, dtype=torch.uint8, device=dev)
    if world > 1:
        dcp.barrier()
    for _ in range(8): flush.zero_()            # ~1.3 ms of queued GPU work: lead time for the host to enqueue three steps
    tr._issue(); tr._issue()
    n1 = len(marks)
    start = torch.cuda.Event(enable_timing=True); start.record()
    tr._issue()
    end = torch.cuda.Event(enable_timing=True); end.record()
    torch.cuda.synchronize()
    K._chk = orig
    rows = []
    prev = {"main": start, "side": None}
    for src, st, e in marks[n1:]:
        t_end = start.elapsed_time(e) * 1e3
        dur = prev[st].elapsed_time(e) * 1e3 if prev[st] is not None else None
        rows.append({"launch": src, "stream": st, "end_us": round(t_end, 2), "since_prev_on_stream_us": None if dur is None else round(dur, 2)})
        prev[st] = e
    total = start.elapsed_time(end) * 1e3
    rec = {"rank": rank, "n_gpus": world, "batch": opt.batch, "total_us": round(total, 2), "launches": rows}
    if world > 1:
        tot = torch.tensor([total], dtype=torch.float64, device=dev)
        allt = [torch.zeros_like(tot) for _ in range(world)]
        dcp.all_gather(allt, tot)
        slow = max(range(world), key=lambda r: allt[r].item())
        rec["totals_us_by_rank"] = [round(t.item(), 2) for t in allt]
        if opt.out and rank == slow and slow != 0:
            json.dump(rec, open(opt.out.replace(".json", f"_rank{rank}_slowest.json"), "w"), indent=1)
    if rank == 0:
        for r in rows:
            d = r["since_prev_on_stream_us"]
            print(f'{r["end_us"]:8.2f} {"" if d is None else f"{d:7.2f}":>7} {r["stream"]:4}  {r["launch"]}')
        print(f"{total:8.2f}  total ({len(rows)} launches)", rec.get("totals_us_by_rank", ""))
        if opt.out:
            json.dump(rec, open(opt.out, "w"), indent=1)
    if world > 1:
        dcp.destroy_process_group()


if __name__ == "__main__":
    main()
