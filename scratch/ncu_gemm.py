"""Short driver for the ncu capture of the persistent bf16 GEMM (lm-head forward shape, 8192 x 50304 x 768)."""
import sys, torch
sys.path.insert(0, '/root/repo')
from distributed_compute_pytorch_b200.ops import cuda as K, tc
K.lib()
M, N, Kd = 8192, 50304, 768
x = torch.randn(M, Kd, device='cuda', dtype=torch.bfloat16); w = torch.randn(N, Kd, device='cuda', dtype=torch.bfloat16) * 0.02
y = torch.empty(M, N, device='cuda', dtype=torch.bfloat16)
for _ in range(3): tc.linear_fwd(x, w, None, y, False)
torch.cuda.synchronize()
print("ok")
