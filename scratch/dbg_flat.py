import sys, ctypes as C, torch
sys.path.insert(0, '/root/repo')
from distributed_compute_pytorch_b200.ops import cuda as K
L = K.lib(); s = K._stream
flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
w = torch.randn(64, 288, device='cuda') * 0.05; b = torch.randn(64, device='cuda')
off = (C.c_int * 9)(*[r * 26 + q for r in range(3) for q in range(3)])
for N in (16, 32, 64, 128, 256, 512, 1024):
    P = N * 676
    a1 = torch.randn(P, 32, device='cuda'); a2 = torch.empty(P, 64, device='cuda')
    for ctas in (148, 74):
        fn = lambda: L.dcp_conv_flat_tc(a1.data_ptr(), w.data_ptr(), a2.data_ptr(), b.data_ptr(), None, P, 32, 64, 9, off, 26, 676, 24, 24, 0, ctas, s())
        for _ in range(3): fn()
        ts = []
        for _ in range(10):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1) * 1e3)
        ts.sort()
        tiles = (P + 127) // 128
        print(f"N={N:5d} tiles={tiles:6d} ctas={ctas} tiles/cta={tiles/ctas:6.1f}  {ts[5]:8.2f} us   per-tile-per-cta {ts[5]/(tiles/ctas):6.2f} us  TF/s {2*P*64*288/ts[5]/1e6:7.1f}")
