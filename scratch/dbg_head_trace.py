"""globaltimer trace of the cluster head kernel (CTA 0, thread 0)."""
import sys, os, torch
sys.path.insert(0, '/root/repo')
dbg = torch.zeros(16, dtype=torch.int64, device='cuda')
os.environ['DCP_HEAD_DBG_PTR'] = str(dbg.data_ptr())
import distributed_compute_pytorch_b200 as dcp
from distributed_compute_pytorch_b200 import ops
from distributed_compute_pytorch_b200.parallel.convnet_trainer import FusedConvNetTrainer
net = dcp.models.ConvNet().cuda().train()
tr = FusedConvNetTrainer(net, ops.Adadelta(net.parameters(), lr=1e-3), 128, use_graph=False)
tr.x.normal_(); tr.y.random_(0, 10)
for _ in range(3): tr._issue()
torch.cuda.synchronize()
flush = torch.empty(1 << 28, dtype=torch.uint8, device='cuda')
names = ["start", "stats done", "bn/relu/dropout done", "partial logits done", "cluster.sync 1 passed", "softmax/dlog done", "fc2 grads + dh3 done", "dh1 written", "end"]
for trial in range(3):
    flush.zero_()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); tr._issue_head_only(); b.record(); torch.cuda.synchronize()
    d = dbg.cpu().tolist()
    print(f"trial {trial}: events {a.elapsed_time(b)*1e3:.1f} us; " + "  ".join(f"{n} +{d[i]-d[0]}" for i, n in enumerate(names)))
