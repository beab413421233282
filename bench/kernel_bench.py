"""Per-kernel device times of the fused ConvNet step (CUDA events, warm-up, L2 flushed between repetitions),
plus achieved FLOP/s / bytes/s against MEASURED_PEAKS.json.  Usage: python bench/kernel_bench.py [--batch 128] [--out file.json]"""
import argparse, ctypes as C, json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import distributed_compute_pytorch_b200 as dcp
from distributed_compute_pytorch_b200 import ops
from distributed_compute_pytorch_b200.ops import cuda as K, tc
from distributed_compute_pytorch_b200.parallel.convnet_trainer import FusedConvNetTrainer


def main():
    ap = argparse.ArgumentParser(); ap.add_argument("--batch", type=int, default=128); ap.add_argument("--out", default=""); ap.add_argument("--reps", type=int, default=20)
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
    rows = []
    for name, (fn, flops, nbytes) in kernels.items():
        for _ in range(3): fn()
        torch.cuda.synchronize()
        cold, warm = [], []
        for mode in ("cold", "warm"):
            for _ in range(opt.reps):
                if mode == "cold":
                    flush.zero_()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record(); fn(); b.record(); torch.cuda.synchronize()
                (cold if mode == "cold" else warm).append(a.elapsed_time(b) * 1e3)
        cold.sort(); warm.sort()
        tc_us, tw_us = cold[len(cold) // 2], warm[len(warm) // 2]
        row = {"kernel": name, "us_L2_flushed": round(tc_us, 2), "us_warm": round(tw_us, 2), "GFLOP": round(flops / 1e9, 3), "MB": round(nbytes / 1e6, 2),
               "TFLOPs_flushed": round(flops / tc_us / 1e6, 2) if flops else None, "GBps_flushed": round(nbytes / tc_us / 1e3, 1),
               "frac_of_measured_hbm": round(nbytes / tc_us / 1e3 / hbm, 3), "frac_of_measured_bf16_gemm_peak": round(flops / tc_us / 1e6 / tfl, 4) if flops else None}
        rows.append(row)
        print(json.dumps(row), flush=True)
    print(json.dumps({"sum_us_L2_flushed": round(sum(r["us_L2_flushed"] for r in rows), 1), "sum_us_warm": round(sum(r["us_warm"] for r in rows), 1)}))
    if opt.out:
        json.dump(rows, open(opt.out, "w"), indent=1)


if __name__ == "__main__":
    main()
