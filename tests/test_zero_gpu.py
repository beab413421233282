"""Fused reduce_scatter -> optimizer -> all_gather step (parallel/zero.py, csrc/coll/fused_opt.cu) against a
single-process PyTorch optimizer fed with the rank-averaged gradients."""
import pytest
import torch
import torch.nn.functional as F
from torch import nn

import distributed_compute_pytorch_b200 as dcp
from distributed_compute_pytorch_b200.parallel.zero import FusedShardedOptimizer
from tests.helpers import free_port, init

pytestmark = [pytest.mark.gpu]


class Net(nn.Module):
    def __init__(self):
        super().__init__()
        self.a = nn.Linear(64, 300); self.b = nn.Linear(300, 77); self.c = nn.Linear(77, 10)
    def forward(self, x):
        return self.c(F.relu(self.b(F.relu(self.a(x)))))


def _worker(rank, world, port, kind, use_nvls):
    torch.cuda.set_device(rank)
    dist = init(rank, world, port, "nccl")
    dev = torch.device("cuda", rank)
    torch.manual_seed(0)
    model = Net().to(dev)
    ref = Net().to(dev); ref.load_state_dict(model.state_dict())
    ddp = dcp.DistributedDataParallel(model, bucket_cap_mb=0.05)
    kw = dict(adam=dict(lr=1e-2, eps=1e-2), adamw=dict(lr=1e-2, weight_decay=0.1, eps=1e-2), sgd=dict(lr=0.1, momentum=0.9), adadelta=dict(lr=1.0))[kind]
    opt = FusedShardedOptimizer(ddp, kind=kind, use_nvls=use_nvls, **kw)
    assert len(ddp._buckets) > 1
    ropt = dict(adam=lambda p: torch.optim.Adam(p, lr=1e-2, eps=1e-2), adamw=lambda p: torch.optim.AdamW(p, lr=1e-2, weight_decay=0.1, eps=1e-2),
                sgd=lambda p: torch.optim.SGD(p, lr=0.1, momentum=0.9), adadelta=lambda p: torch.optim.Adadelta(p, lr=1.0))[kind](ref.parameters())
    g = torch.Generator().manual_seed(5)
    X, Y = torch.randn(5, 8 * world, 64, generator=g).to(dev), torch.randint(0, 10, (5, 8 * world), generator=g).to(dev)
    for step in range(5):
        opt.zero_grad()
        F.cross_entropy(ddp(X[step, rank * 8:(rank + 1) * 8]), Y[step, rank * 8:(rank + 1) * 8]).backward()
        opt.step()
        ropt.zero_grad()
        F.cross_entropy(ref(X[step]), Y[step]).backward()
        ropt.step()
        torch.cuda.synchronize()
        for (n, p), q in zip(model.named_parameters(), ref.parameters()):
            assert torch.allclose(p, q, rtol=2e-4, atol=2e-5), (kind, step, n, (p - q).abs().max().item())
    flat = torch.cat([p.detach().reshape(-1) for p in model.parameters()])
    both = [torch.empty_like(flat) for _ in range(world)]
    dist.all_gather(both, flat)
    assert all(torch.equal(both[0], b) for b in both)          # replicas bit-identical
    assert all(k.startswith("module.") for k in ddp.state_dict())
    assert tuple(ddp.state_dict()["module.a.weight"].shape) == (300, 64)
    dist.destroy_process_group()


@pytest.mark.multigpu
@pytest.mark.parametrize("kind,use_nvls", [("adam", True), ("adamw", False), ("sgd", True), ("adadelta", False)])
def test_fused_sharded_optimizer_matches_torch(kind, use_nvls):
    world = min(torch.cuda.device_count(), 4)
    dcp.spawn(_worker, args=(world, free_port(), kind, use_nvls), nprocs=world, join=True)
