"""Causal attention forward / backward: the tcgen05 kernels of csrc/ops/attention_tc.cu against torch SDPA (cuDNN / flash).

GPT-2 small shape by default (B=8, T=1024, H=12, D=64, bf16).  Each measurement is the median of `--reps` CUDA-event timings
with an L2 flush (a 256 MiB write) between repetitions; `--once` runs one forward + backward of ours only (for ncu).

    python bench/attn_bench.py [--B 8 --T 1024 --H 12] [--out profiles/attn_bench.json]
"""
import argparse
import json
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_compute_pytorch_b200.ops import cuda as K  # noqa: E402


def timed(fn, reps, flush):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(reps):
        flush.fill_(1.0)
        torch.cuda._sleep(600000)      # let the host run ahead: the events then bracket device time only
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--B", type=int, default=8); ap.add_argument("--T", type=int, default=1024); ap.add_argument("--H", type=int, default=12)
    ap.add_argument("--reps", type=int, default=30)
    ap.add_argument("--once", action="store_true")
    ap.add_argument("--out", default="")
    ap.add_argument("--trace-cta", default="0,0", help="bx,which: CTA index; which = 0 backward, 1 forward")
    ap.add_argument("--trace", action="store_true", help="clock64 stamps of backward CTA (0, 0): per query block, cycles since the first stamp")
    a = ap.parse_args()
    B, T, H, D = a.B, a.T, a.H, 64
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    qkv = torch.randn(B * T, 3 * H * D, device=dev, dtype=torch.bfloat16)
    dout = torch.randn(B * T, H * D, device=dev, dtype=torch.bfloat16)
    if a.trace:
        import ctypes
        lib = K.lib()
        lib.dcp_attn_set_trace.argtypes, lib.dcp_attn_set_trace.restype = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int], None
        buf = torch.zeros(8 * 32, device=dev, dtype=torch.int64)
        x = qkv.clone().requires_grad_(True)
        y = K.causal_attention(x, B, T, H)
        y.backward(dout)
        lib.dcp_attn_set_trace(buf.data_ptr(), *map(int, a.trace_cta.split(',')))
        x.grad = None
        K.causal_attention(x, B, T, H).backward(dout)
        torch.cuda.synchronize()
        lib.dcp_attn_set_trace(None, 0, 0)
        st = buf.cpu().tolist()
        t0 = min(v for v in st if v)
        if a.trace_cta.endswith(",1"):
            print("forward stamps per key block: thr s_full, pass1 done, o updated, P stored; mma p_ready, S(j+1) issued, PV issued")
        names = ["thr: S/dP ready", "thr: P/dS in regs", "thr: dQ(it-1) taken", "thr: P/dS stored", "mma: P/dS ready", "mma: S/dP(it+1) issued", "mma: dV/dK/dQ issued", "thr: dQ complete"]
        for it in range(32):
            print(it, {n: st[it * 8 + k] - t0 for k, n in enumerate(names) if st[it * 8 + k]})
        return
    if a.once:
        x = qkv.clone().requires_grad_(True)
        K.causal_attention(x, B, T, H).backward(dout)
        torch.cuda.synchronize()
        print("ok")
        return
    flush = torch.empty(64 << 20, device=dev)
    # ours
    x = qkv.clone().requires_grad_(True)
    y = K.causal_attention(x, B, T, H)
    ours_f = timed(lambda: K.causal_attention(qkv, B, T, H), a.reps, flush)
    ours_b = timed(lambda: torch.autograd.grad(y, x, dout, retain_graph=True), a.reps, flush)
    # library: SDPA on [B, H, T, D] views of the same matrix
    xs = qkv.clone().requires_grad_(True)
    q, k, v = (t.view(B, T, H, D).transpose(1, 2) for t in xs.view(B, T, 3 * H * D).split(H * D, dim=2))
    do = dout.view(B, T, H, D).transpose(1, 2)
    res = {"B": B, "T": T, "H": H, "D": D, "ours_fwd_us": ours_f, "ours_bwd_us": ours_b}
    from torch.nn.attention import SDPBackend, sdpa_kernel
    for name, be in (("cudnn", SDPBackend.CUDNN_ATTENTION), ("flash", SDPBackend.FLASH_ATTENTION)):
        try:
            with sdpa_kernel(be):
                yl = F.scaled_dot_product_attention(q, k, v, is_causal=True)
                res[f"{name}_fwd_us"] = timed(lambda: F.scaled_dot_product_attention(q, k, v, is_causal=True), a.reps, flush)
                res[f"{name}_bwd_us"] = timed(lambda: torch.autograd.grad(yl, (q, k, v), do, retain_graph=True), a.reps, flush)
        except Exception as e:  # backend not available for this shape
            res[f"{name}_error"] = str(e)[:120]
    flops = 4.0 * B * H * T * T * D / 2
    res["ours_fwd_tflops"] = flops / ours_f / 1e6
    res["ours_bwd_tflops"] = 2.5 * flops / ours_b / 1e6
    line = json.dumps(res)
    print(line)
    if a.out:
        with open(a.out, "w") as f:
            f.write(line + "\n")


if __name__ == "__main__":
    main()
