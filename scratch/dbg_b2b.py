import sys, ctypes as C, torch
sys.path.insert(0, '/root/repo')
from distributed_compute_pytorch_b200.ops import cuda as K
L = K.lib(); s = K._stream
flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
w = torch.randn(64, 288, device='cuda') * 0.05; b = torch.randn(64, device='cuda')
off = (C.c_int * 9)(*[r * 26 + q for r in range(3) for q in range(3)])
N = 128; P = N * 676
a1 = torch.randn(P, 32, device='cuda'); a2 = torch.empty(P, 64, device='cuda')
small = torch.randn(1 << 20, device='cuda')
fn = lambda: L.dcp_conv_flat_tc(a1.data_ptr(), w.data_ptr(), a2.data_ptr(), b.data_ptr(), None, P, 32, 64, 9, off, 26, 676, 24, 24, 0, 148, s())
for _ in range(3): fn()
def seq(pre, n=4):
    res = []
    for rep in range(5):
        pre()
        evs = [torch.cuda.Event(enable_timing=True) for _ in range(n + 1)]
        evs[0].record()
        for i in range(n):
            fn(); evs[i + 1].record()
        torch.cuda.synchronize()
        res.append([evs[i].elapsed_time(evs[i + 1]) * 1e3 for i in range(n)])
    return [sorted(x)[2] for x in zip(*res)]
print('after 256MB flush   :', [round(v, 1) for v in seq(lambda: flush.zero_())])
print('after small torch op:', [round(v, 1) for v in seq(lambda: small.mul_(1.0001))])
print('after nothing (sync):', [round(v, 1) for v in seq(lambda: torch.cuda.synchronize())])
g = torch.cuda.CUDAGraph()
st = torch.cuda.Stream()
with torch.cuda.stream(st):
    fn(); torch.cuda.synchronize()
    with torch.cuda.graph(g):
        for _ in range(10): fn()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
g.replay(); torch.cuda.synchronize()
e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
print('graph of 10 back-to-back convs: per kernel', round(e0.elapsed_time(e1) * 100, 1), 'us')
