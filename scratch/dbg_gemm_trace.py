"""clock64 trace of CTA (0,0,0) of the three fc1 GEMMs (3xTF32) - where do 20 us go for 4 K stages?"""
import sys, os, torch
sys.path.insert(0, '/root/repo')
dbg = torch.zeros(128, dtype=torch.int64, device='cuda')
os.environ['DCP_GEMM_DBG_PTR'] = str(dbg.data_ptr())
from distributed_compute_pytorch_b200.ops import cuda as K, tc
K.lib()
N = 128
pooled = torch.randn(N, 9216, device='cuda'); W = torch.randn(128, 9216, device='cuda') * 0.01; b = torch.randn(128, device='cuda')
h1 = torch.empty(N, 128, device='cuda'); dh1 = torch.randn(N, 128, device='cuda'); dW = torch.empty(128, 9216, device='cuda'); dp = torch.empty(N, 9216, device='cuda')
flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
cases = {"fc1_fwd": lambda: tc.linear_fwd(pooled, W, b, h1, False),
         "fc1_wgrad": lambda: tc.gemm(dh1, pooled, dW, None, 128, 9216, N, 128, 9216, 9216, 1, 1, False, 1),
         "fc1_dgrad": lambda: tc.gemm(dh1, W, dp, None, N, 9216, 128, 128, 9216, 9216, 0, 1, False, 1)}
for x3 in (1, 0):
    tc.LINEAR_X3 = bool(x3)
    for name, fn in cases.items():
        for cold in (0, 1):
            for _ in range(3): fn()
            torch.cuda.synchronize()
            if cold: flush.zero_()
            dbg.zero_(); torch.cuda.synchronize()
            a, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); fn(); e.record(); torch.cuda.synchronize()
            d = dbg.cpu().tolist(); t0 = d[0]
            print(f"{name} x3={x3} cold={cold}: {a.elapsed_time(e)*1e3:.1f} us | setup +{d[1]-t0} acc_full seen +{d[3]-t0} epi done +{d[4]-t0} end +{d[2]-t0} (cycles)")
            for i in range(4):
                r = [x - t0 if x else 0 for x in d[8+i*4: 12+i*4]]
                print(f"   stage {i}: tma issued +{r[0]}  split-warp full +{r[3]}  mma-warp ready +{r[1]}  mma issued +{r[2]}")
