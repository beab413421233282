import sys, os, ctypes as C, torch
sys.path.insert(0, '/root/repo')
dbg = torch.zeros(128, dtype=torch.int64, device='cuda')
os.environ['DCP_FLAT_DBG_PTR'] = str(dbg.data_ptr())
from distributed_compute_pytorch_b200.ops import cuda as K
L = K.lib(); s = K._stream
w = torch.randn(64, 288, device='cuda') * 0.05; b = torch.randn(64, device='cuda')
off = (C.c_int * 9)(*[r * 26 + q for r in range(3) for q in range(3)])
N = 256; P = N * 676
a1 = torch.randn(P, 32, device='cuda'); a2 = torch.empty(P, 64, device='cuda')
for _ in range(3):
    L.dcp_conv_flat_tc(a1.data_ptr(), w.data_ptr(), a2.data_ptr(), b.data_ptr(), None, P, 32, 64, 9, off, 26, 676, 24, 24, 0, 148, s())
torch.cuda.synchronize()
d = dbg.cpu().tolist(); t0 = d[0]
print('weights ready +', d[1]-t0, ' kernel end +', d[2]-t0)
for i in range(8):
    r = [x - t0 for x in d[8+i*8: 8+i*8+8]]
    print(f'tile {i}: full ok +{r[1]}  mma issued +{r[2]}  epi wake +{r[3]}  c0:tmem_ld done +{r[5]}  c0:sts done +{r[6]}  c0:stg done +{r[7]}  epi done +{r[4]}')
