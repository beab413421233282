"""bf16 GEMM throughput on the GPT-2-small / ResNet-50 Linear shapes: this package's tcgen05 kernels (persistent 128x256
kernel; DCP_GEMM_PERSISTENT=0 selects the older one-tile-per-CTA kernel) against cuBLAS (torch.matmul) on the same box.
CUDA events, warm-up, operands larger than ... rotated between repetitions (4 operand sets) so L2 does not flatter either.

    python bench/gemm_bench.py [--out profiles/gemm_bench.json]
"""
import argparse, json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_compute_pytorch_b200.ops import cuda as K, tc

T = 8192   # tokens per GPU (batch 8 x seq 1024)
SHAPES = [  # (name, kind, M, N, K)   fwd: y[M,N] = x[M,K] w[N,K]^T ; dgrad: dx[M,K'] ; wgrad: dw
    ("qkv fwd", "fwd", T, 2304, 768), ("attn-proj fwd", "fwd", T, 768, 768), ("mlp-fc fwd", "fwd", T, 3072, 768), ("mlp-proj fwd", "fwd", T, 768, 3072),
    ("lm-head fwd", "fwd", T, 50304, 768),
    ("mlp-fc dgrad", "dgrad", T, 3072, 768), ("lm-head dgrad", "dgrad", T, 50304, 768),
    ("mlp-fc wgrad", "wgrad", T, 3072, 768), ("attn-proj wgrad", "wgrad", T, 768, 768), ("lm-head wgrad", "wgrad", T, 50304, 768),
    ("square 8192", "fwd", 8192, 8192, 8192),
]


def timed(fn, sets, reps):
    for i in range(3): fn(i % sets)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(reps): fn(i % sets)
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps * 1e-3


def main():
    ap = argparse.ArgumentParser(); ap.add_argument("--out", default=""); ap.add_argument("--reps", type=int, default=20)
    opt = ap.parse_args()
    K.lib()
    dev = "cuda"
    rows = []
    for name, kind, M, N, Kd in SHAPES:
        sets = 2 if M * N > 1 << 28 else 4
        x = [torch.randn(M, Kd, device=dev, dtype=torch.bfloat16) for _ in range(sets)]
        w = [torch.randn(N, Kd, device=dev, dtype=torch.bfloat16) * 0.02 for _ in range(sets)]
        flops = 2.0 * M * N * Kd
        if kind == "fwd":
            y = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
            ours = lambda i: tc.linear_fwd(x[i], w[i], None, y, False)
            ref = lambda i: torch.matmul(x[i], w[i].t(), out=y)
        elif kind == "dgrad":
            dy = [torch.randn(M, N, device=dev, dtype=torch.bfloat16) for _ in range(sets)]
            dx = torch.empty(M, Kd, device=dev, dtype=torch.bfloat16)
            ours = lambda i: tc.linear_dgrad(dy[i], w[i], dx)
            ref = lambda i: torch.matmul(dy[i], w[i], out=dx)
        else:
            dy = [torch.randn(M, N, device=dev, dtype=torch.bfloat16) for _ in range(sets)]
            dw32 = torch.empty(N, Kd, device=dev, dtype=torch.float32); dwb = torch.empty(N, Kd, device=dev, dtype=torch.bfloat16)
            ours = lambda i: tc.linear_wgrad(dy[i], x[i], dw32, None)
            ref = lambda i: torch.matmul(dy[i].t(), x[i], out=dwb)
        t_o, t_r = timed(ours, sets, opt.reps), timed(ref, sets, opt.reps)
        row = {"gemm": name, "M": M, "N": N, "K": Kd, "ours_us": round(t_o * 1e6, 1), "cublas_us": round(t_r * 1e6, 1),
               "ours_TFLOPs": round(flops / t_o / 1e12, 1), "cublas_TFLOPs": round(flops / t_r / 1e12, 1), "ours_vs_cublas": round(t_r / t_o, 3)}
        rows.append(row); print(json.dumps(row), flush=True)
        del x, w
    if opt.out:
        json.dump(rows, open(opt.out, "w"), indent=1)


if __name__ == "__main__":
    main()
