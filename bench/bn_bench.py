"""Isolated timing of the BatchNorm kernels (dcp_bn_stats / apply / bwd) on ResNet-50 shaped activations."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_compute_pytorch_b200.ops import cuda as K
L = K.lib(); s = K._stream
def t(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3
for (N, C, H) in [(64, 64, 112), (64, 256, 56), (64, 512, 28), (64, 2048, 7)]:
    M = N * H * H
    xs = [torch.randn(M, C, device='cuda').to(torch.bfloat16) for _ in range(3)]
    y = torch.empty_like(xs[0]); dy = torch.randn(M, C, device='cuda').to(torch.bfloat16); dx = torch.empty_like(y)
    g = torch.ones(C, device='cuda'); b = torch.zeros(C, device='cuda'); rm = torch.zeros(C, device='cuda'); rv = torch.ones(C, device='cuda')
    nbt = torch.zeros((), dtype=torch.int64, device='cuda'); sm = torch.empty(C, device='cuda'); sr = torch.empty(C, device='cuda')
    scr = torch.zeros(4 * C + 2, dtype=torch.float64, device='cuda'); dg = torch.empty(C, device='cuda'); db = torch.empty(C, device='cuda')
    i = [0]
    def nx(): i[0] += 1; return xs[i[0] % 3]
    mb = M * C * 2 / 1e6
    r = {}
    r['stats'] = (t(lambda: L.dcp_bn_stats(nx().data_ptr(), scr.data_ptr(), M, C, 1, s())), mb)
    r['apply'] = (t(lambda: L.dcp_bn_apply(nx().data_ptr(), y.data_ptr(), scr.data_ptr(), g.data_ptr(), b.data_ptr(), rm.data_ptr(), rv.data_ptr(), nbt.data_ptr(), sm.data_ptr(), sr.data_ptr(), M, C, float(M), 1e-5, 0.1, 1, 1, 1, s())), 2 * mb)
    r['bwd_reduce'] = (t(lambda: L.dcp_bn_bwd_reduce(nx().data_ptr(), y.data_ptr(), dy.data_ptr(), sm.data_ptr(), sr.data_ptr(), scr.data_ptr(), M, C, 1, 1, s())), 3 * mb)
    r['bwd_apply'] = (t(lambda: L.dcp_bn_bwd_apply(nx().data_ptr(), y.data_ptr(), dy.data_ptr(), dx.data_ptr(), sm.data_ptr(), sr.data_ptr(), g.data_ptr(), scr.data_ptr(), dg.data_ptr(), db.data_ptr(), M, C, float(M), 1, 1, s())), 4 * mb)
    print(f"M={M} C={C} ({mb:.0f} MB): " + "  ".join(f"{k} {v[0]:.1f}us {v[1]/v[0]*1e-3*1e3:.0f}GB/s" for k, v in r.items()), flush=True)
