"""Back-to-back timeline of ONE fused ConvNet training step: every launch of FusedConvNetTrainer._issue() is followed
by a CUDA event, the whole step is queued behind a long memset so the host never starves the GPU, and the gaps between
consecutive events give each kernel's duration under the conditions of the real (graph-replayed) step - warm L2,
predecessor just finished.  The second of two queued steps is reported.  (The event records themselves cost a little;
the graph-replayed step of bench.py is the authoritative total.)

    python bench/step_timeline.py [--batch 128] [--out profiles/step_timeline.json]
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
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    net = dcp.models.ConvNet().to(dev).train()
    optim = ops.Adadelta(net.parameters(), lr=1e-3)
    tr = FusedConvNetTrainer(net, optim, opt.batch, use_graph=False)
    tr.x.normal_(); tr.y.random_(0, 10)
    for _ in range(3): tr._issue()
    torch.cuda.synchronize()
    marks = []
    orig = K._chk

    def chk(rc):
        orig(rc)
        f = g = sys._getframe(1)
        while g is not None and not g.f_code.co_filename.endswith("convnet_trainer.py"):
            g = g.f_back
        f = g or f
        src = linecache.getline(f.f_code.co_filename, f.f_lineno).strip()
        e = torch.cuda.Event(enable_timing=True); e.record()
        marks.append((src[:90], e))
    K._chk = chk
    flush = torch.empty(1 << 30, dtype=torch.uint8, device=dev)
    for _ in range(6): flush.zero_()            # ~1 ms of queued GPU work: lead time for the host to enqueue two steps
    tr._issue()
    n1 = len(marks)
    start = torch.cuda.Event(enable_timing=True); start.record()
    tr._issue()
    torch.cuda.synchronize()
    K._chk = orig
    rows, prev = [], start
    for src, e in marks[n1:]:
        rows.append({"launch": src, "us": round(prev.elapsed_time(e) * 1e3, 2)}); prev = e
    for r in rows: print(f'{r["us"]:8.2f}  {r["launch"]}')
    total = start.elapsed_time(marks[-1][1]) * 1e3
    print(f"{total:8.2f}  total ({len(rows)} launches)")
    if opt.out:
        json.dump({"batch": opt.batch, "total_us": total, "launches": rows}, open(opt.out, "w"), indent=1)


if __name__ == "__main__":
    main()
