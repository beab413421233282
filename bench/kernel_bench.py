"""Per-kernel device times of the fused ConvNet step (CUDA events, warm-up, L2 flushed between repetitions),
plus achieved FLOP/s / bytes/s against MEASURED_PEAKS.json.

--with-library adds, for every kernel, the library counterpart the reference model runs through (cuDNN convolutions with
allow_tf32 as torch ships it, NCHW and channels-last - the better of the two is kept; cuBLAS fp32 and tf32 GEMMs for fc1's three
shapes; ATen max-pool / dropout / BatchNorm / log-softmax / nll; torch.optim.Adadelta foreach) on the same box, same timing:
each callable is warmed up, captured ONCE in a CUDA graph (so a multi-kernel library sequence is not charged for the host's
launch gaps) and the replay is timed with CUDA events - L2 flushed before every replay ("cold"), back to back ("warm").

    python bench/kernel_bench.py [--batch 128] [--with-library] [--out file.json]"""
import argparse, ctypes as C, json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import distributed_compute_pytorch_b200 as dcp
from distributed_compute_pytorch_b200 import ops
from distributed_compute_pytorch_b200.ops import cuda as K, tc
from distributed_compute_pytorch_b200.parallel.convnet_trainer import FusedConvNetTrainer


def library_rows(N, dev, measure):
    """The PyTorch-library counterpart of every kernel above (what the reference model executes, ref:main.py:32-44,61-63), timed the same way."""
    import torch.nn.functional as F
    rows = []

    def add(name, ours, fn):
        c, w = measure(fn)
        row = {"kernel": "LIB " + name, "counterpart_of": ours, "us_L2_flushed": round(c, 2), "us_warm": round(w, 2)}
        rows.append(row); print(json.dumps(row), flush=True)

    f32 = dict(device=dev, dtype=torch.float32)
    cb = torch.ops.aten.convolution_backward
    for fmt_name, fmt in (("NCHW", torch.contiguous_format), ("channels_last", torch.channels_last)):
        x1 = torch.randn(N, 1, 28, 28, **f32).contiguous(memory_format=fmt); w1 = torch.randn(32, 1, 3, 3, **f32).contiguous(memory_format=fmt); b1 = torch.randn(32, **f32)
        a1 = torch.randn(N, 32, 26, 26, **f32).contiguous(memory_format=fmt); w2 = torch.randn(64, 32, 3, 3, **f32).contiguous(memory_format=fmt); b2 = torch.randn(64, **f32)
        dy2 = torch.randn(N, 64, 24, 24, **f32).contiguous(memory_format=fmt); dz1 = torch.randn(N, 32, 26, 26, **f32).contiguous(memory_format=fmt)
        a2 = torch.randn(N, 64, 24, 24, **f32).contiguous(memory_format=fmt)
        add(f"cuDNN conv1 fwd + relu ({fmt_name}, tf32 allowed)", "conv1_fwd (SIMT)", lambda: F.relu(F.conv2d(x1, w1, b1)))
        add(f"cuDNN conv2 fwd (bias, no relu) ({fmt_name})", "conv2_fwd flat tcgen05 (tf32)", lambda: F.conv2d(a1, w2, b2))
        add(f"cuDNN conv2 fwd + relu ({fmt_name})", "conv2_fwd flat tcgen05 (tf32)", lambda: F.relu(F.conv2d(a1, w2, b2)))
        add(f"cuDNN conv2 dgrad ({fmt_name})", "conv2_dgrad flat tcgen05 (tf32)",
            lambda: cb(dy2, a1, w2, [64], [1, 1], [0, 0], [1, 1], False, [0, 0], 1, [True, False, False]))
        add(f"cuDNN conv2 wgrad + bias grad ({fmt_name})", "conv2_wgrad+bias_grad FLAT tcgen05 (tf32) + reduce",
            lambda: cb(dy2, a1, w2, [64], [1, 1], [0, 0], [1, 1], False, [0, 0], 1, [False, True, True]))
        add(f"cuDNN conv1 wgrad + bias grad ({fmt_name})", "conv1_wgrad (SIMT)",
            lambda: cb(dz1, x1, w1, [32], [1, 1], [0, 0], [1, 1], False, [0, 0], 1, [False, True, True]))
        add(f"ATen max_pool2d + dropout2d fwd ({fmt_name})", "pool_fwd+dropout", lambda: F.dropout2d(F.max_pool2d(F.relu(a2), 2), 0.25, True))
        rl = F.relu(a2)
        pooled, idx = F.max_pool2d(rl, 2, return_indices=True)
        mask = (torch.rand(N, 64, 1, 1, device=dev) >= 0.25).float() / 0.75
        gp = torch.randn_like(pooled)
        add(f"ATen dropout2d + max_pool2d + relu bwd ({fmt_name})", "pool_bwd (+relu/dropout bwd)",
            lambda: torch.ops.aten.threshold_backward(
                torch.ops.aten.max_pool2d_with_indices_backward(gp * mask, rl, [2, 2], [2, 2], [0, 0], [1, 1], False, idx), rl, 0))
    xf = torch.randn(N, 9216, **f32); wf = torch.randn(128, 9216, **f32); bf = torch.randn(128, **f32); dh = torch.randn(N, 128, **f32)
    for tf32 in (False, True):
        torch.backends.cuda.matmul.allow_tf32 = tf32
        tag = "tf32" if tf32 else "fp32 (torch default, what the reference gets)"
        add(f"cuBLAS fc1 fwd {tag}", "fc1_fwd tcgen05 3xTF32 split-K+reduce", lambda: F.linear(xf, wf, bf))
        add(f"cuBLAS fc1 wgrad {tag}", "fc1_wgrad tcgen05 3xTF32", lambda: dh.t() @ xf)
        add(f"cuBLAS fc1 dgrad {tag}", "fc1_dgrad tcgen05 3xTF32", lambda: dh @ wf)
    torch.backends.cuda.matmul.allow_tf32 = False
    # head: BatchNorm1d -> relu -> dropout -> fc2 -> log_softmax -> nll, forward AND backward (autograd), as the reference runs it
    bn = torch.nn.BatchNorm1d(128).to(dev).train(); fc2 = torch.nn.Linear(128, 10).to(dev)
    h1 = torch.randn(N, 128, **f32, requires_grad=True); y = torch.randint(0, 10, (N,), device=dev)
    params = [h1] + list(bn.parameters()) + list(fc2.parameters())

    def head():
        out = F.log_softmax(fc2(F.dropout(F.relu(bn(h1)), 0.5, True)), dim=1)
        return torch.autograd.grad(F.nll_loss(out, y), params)
    add("ATen/cuBLAS head: BatchNorm1d..nll_loss forward + backward (autograd, ~30 kernels)", "head megakernel (BN..CE fwd+bwd)", head)
    net = dcp.models.ConvNet().to(dev)
    for q in net.parameters():
        q.grad = torch.randn_like(q)
    topt = torch.optim.Adadelta(net.parameters(), lr=1e-3, foreach=True, capturable=True)
    add("torch.optim.Adadelta foreach (capturable)", "adadelta (multi-tensor)", lambda: topt.step())
    return rows


def main():
    ap = argparse.ArgumentParser(); ap.add_argument("--batch", type=int, default=128); ap.add_argument("--out", default=""); ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--with-library", action="store_true")
    opt = ap.parse_args()
    N = opt.batch
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    net = dcp.models.ConvNet().to(dev).train()
    optim = ops.Adadelta(net.parameters(), lr=1e-3)
    tr = FusedConvNetTrainer(net, optim, N, use_graph=False)
    tr.x.normal_(); tr.y.random_(0, 10)
    for _ in range(3): tr._issue()
    torch.cuda.synchronize()
    L, s = K.lib(), K._stream
    p = dict(net.named_parameters()); g = {n: t.grad for n, t in p.items()}
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    P = N * 676
    step_ptr = K.graph_step.data_ptr()
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm = peaks.get("hbm_gbs", 6650.0); tfl = peaks.get("bf16_tflops", 1590.0)
    kernels = {
        "conv1_fwd (SIMT)": (lambda: L.dcp_convnet_conv1_fwd(tr.x.data_ptr(), p["conv1.weight"].data_ptr(), p["conv1.bias"].data_ptr(), tr.a1.data_ptr(), N, s()),
                             2 * P * 32 * 9, N * 784 * 4 + P * 32 * 4),
        "conv2_fwd flat tcgen05 (tf32)": (lambda: L.dcp_conv_flat_tc(tr.a1.data_ptr(), p["conv2.weight"].data_ptr(), tr.a2.data_ptr(), p["conv2.bias"].data_ptr(), None, P, 32, 64, 9, tr.off_fwd, 26, 676, 24, 24, 0, 148, s()),
                                          2 * P * 64 * 288, P * 32 * 4 + P * 64 * 4),
        "pool_fwd+dropout": (lambda: L.dcp_convnet_pool_fwd(tr.a2.data_ptr(), tr.pooled.data_ptr(), tr.arg.data_ptr(), N, 0.25, tr.seed, 1, step_ptr, 1, s()), 0, P * 64 * 4 + N * 9216 * 5),
        "fc1_fwd tcgen05 3xTF32 split-K+reduce": (lambda: tc.linear_fwd(tr.pooled, p["fc1.weight"], p["fc1.bias"], tr.h1, False), 3 * 2 * N * 128 * 9216, (N + 128) * 9216 * 4),
        "head megakernel (BN..CE fwd+bwd)": (lambda: tr._issue_head_only(), 0, N * 128 * 8),
        "fc1_wgrad tcgen05 3xTF32": (lambda: tc.gemm(tr.dh1, tr.pooled, g["fc1.weight"], None, 128, 9216, N, 128, 9216, 9216, 1, 1, False, 1), 3 * 2 * N * 128 * 9216, (N + 128) * 9216 * 4),
        "fc1_dgrad tcgen05 3xTF32": (lambda: tc.gemm(tr.dh1, p["fc1.weight"], tr.dpooled, None, N, 9216, 128, 128, 9216, 9216, 0, 1, False, 1), 3 * 2 * N * 128 * 9216, (N + 128) * 9216 * 4),
        "pool_bwd (+relu/dropout bwd)": (lambda: L.dcp_convnet_pool_bwd(tr.dpooled.data_ptr(), tr.pooled.data_ptr(), tr.arg.data_ptr(), tr.dy2.data_ptr(), N, 0.25, tr.seed, 1, step_ptr, 1, s()), 0, P * 64 * 4 + N * 9216 * 9),
        "conv2_wgrad tcgen05 (tf32) + reduce": (lambda: tc.conv_wgrad_launch(tr.dy2, tr.a1, g["conv2.weight"], N, 26, 26, 32, 64, 3, 3, 0, 26, 26, 8, 4, 1, 148), 2 * P * 64 * 288, P * 96 * 4),
        "conv2_wgrad+bias_grad FLAT tcgen05 (tf32) + reduce": (lambda: (L.dcp_conv_wgrad_flat_tc(tr.a1.data_ptr(), tr.dy2.data_ptr(), tr.wg_scratch.data_ptr(), P, 64, 26, tr.wg_parts, s()),
                                                              tc._lib.dcp_splitk_reduce2(tr.wg_scratch.data_ptr(), g["conv2.weight"].data_ptr(), g["conv2.bias"].data_ptr(), 64 * 288, tr.wg_parts, 64 * 289, s())), 2 * P * 64 * 288, P * 96 * 4),
        "conv2_bias_grad colsum": (lambda: L.dcp_colsum(tr.dy2.data_ptr(), g["conv2.bias"].data_ptr(), P, 64, 0, s()), 0, P * 64 * 4),
        "conv2_dgrad flat tcgen05 (tf32)": (lambda: L.dcp_conv_flat_tc(tr.dy2.data_ptr(), tr.wt2.data_ptr(), tr.dz1.data_ptr(), None, tr.a1.data_ptr(), P, 64, 32, 9, tr.off_bwd, 26, 676, 26, 26, 1, 148, s()),
                                            2 * P * 32 * 576, P * 64 * 4 + P * 32 * 8),
        "conv1_wgrad (SIMT)": (lambda: L.dcp_convnet_conv1_wgrad(tr.dz1.data_ptr(), tr.x.data_ptr(), g["conv1.weight"].data_ptr(), g["conv1.bias"].data_ptr(), N, s()), 2 * P * 32 * 10, P * 32 * 4),
        "adadelta (multi-tensor)": (lambda: optim.step(), 0, 1200138 * 4 * 7),
    }
    from distributed_compute_pytorch_b200.utils import ClockSampler

    def measure(fn):
        """(cold us, warm us): median replay time of fn captured in a CUDA graph, L2 flushed before each replay / back to back."""
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(3): fn()
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr):
            fn()
        gr.replay(); torch.cuda.synchronize()
        cold, warm = [], []
        for mode in ("cold", "warm"):
            for _ in range(opt.reps):
                if mode == "cold":
                    flush.zero_()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record(); gr.replay(); b.record(); torch.cuda.synchronize()
                (cold if mode == "cold" else warm).append(a.elapsed_time(b) * 1e3)
        cold.sort(); warm.sort()
        return cold[len(cold) // 2], warm[len(warm) // 2]

    rows = []
    clocks = ClockSampler(0, period_s=0.05)
    clocks.__enter__()
    for name, (fn, flops, nbytes) in kernels.items():
        tc_us, tw_us = measure(fn)
        row = {"kernel": name, "us_L2_flushed": round(tc_us, 2), "us_warm": round(tw_us, 2), "GFLOP": round(flops / 1e9, 3), "MB": round(nbytes / 1e6, 2),
               "TFLOPs_flushed": round(flops / tc_us / 1e6, 2) if flops else None, "GBps_flushed": round(nbytes / tc_us / 1e3, 1),
               "frac_of_measured_hbm": round(nbytes / tc_us / 1e3 / hbm, 3), "frac_of_measured_bf16_gemm_peak": round(flops / tc_us / 1e6 / tfl, 4) if flops else None}
        rows.append(row)
        print(json.dumps(row), flush=True)
    if opt.with_library:
        rows += library_rows(N, dev, measure)
    clocks.__exit__(None, None, None)
    rows.append({"clocks": clocks.summary()})
    print(json.dumps(rows[-1]), flush=True)
    own = [r for r in rows if "kernel" in r and not r["kernel"].startswith("LIB ")]
    print(json.dumps({"sum_us_L2_flushed": round(sum(r["us_L2_flushed"] for r in own), 1), "sum_us_warm": round(sum(r["us_warm"] for r in own), 1)}))
    if opt.out:
        json.dump(rows, open(opt.out, "w"), indent=1)


if __name__ == "__main__":
    main()
