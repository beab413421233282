"""DDP wrapper + native reducer on the CPU mesh (world_size 2): bucket layout, gradient averaging,
replica consistency, checkpoint key compatibility, no_sync."""
import os
import tempfile

import pytest
import torch
import torch.distributed as tdist
import torch.nn.functional as F
from torch import nn

import distributed_compute_pytorch_b200 as dcp
from distributed_compute_pytorch_b200.models import ConvNet
from distributed_compute_pytorch_b200.parallel.ddp import plan_buckets
from tests.helpers import free_port, init

REF_KEYS = {
    "conv1.weight": (32, 1, 3, 3), "conv1.bias": (32,), "conv2.weight": (64, 32, 3, 3), "conv2.bias": (64,),
    "fc1.weight": (128, 9216), "fc1.bias": (128,), "fc2.weight": (10, 128), "fc2.bias": (10,),
    "batchnorm.weight": (128,), "batchnorm.bias": (128,), "batchnorm.running_mean": (128,),
    "batchnorm.running_var": (128,), "batchnorm.num_batches_tracked": (),
}


def test_bucket_plan_matches_library_for_convnet():
    m = ConvNet()
    names = [n for n, _ in m.named_parameters()]
    ready = ["fc2.bias", "fc2.weight", "batchnorm.weight", "batchnorm.bias", "fc1.bias", "fc1.weight",
             "conv2.weight", "conv2.bias", "conv1.weight", "conv1.bias"]
    params = dict(m.named_parameters())
    nbytes = [params[n].numel() * 4 for n in ready]
    assign = plan_buckets(nbytes, [0] * len(ready), [1024 * 1024, 25 * 1024 * 1024])
    sizes = [sum(b for b, a in zip(nbytes, assign) if a == k) for k in range(max(assign) + 1)]
    assert sizes == [4725288, 75264]
    # same answer from the library planner
    lib_buckets, _ = tdist._compute_bucket_assignment_by_size([params[n] for n in ready], [1024 * 1024, 25 * 1024 * 1024])
    lib_assign = [None] * len(ready)
    for bi, idxs in enumerate(lib_buckets):
        for i in idxs:
            lib_assign[i] = bi
    assert assign == lib_assign
    assert sum(nbytes) == 4800552 and len(names) == 10
    # one unbounded bucket for iteration 1
    assert set(plan_buckets(nbytes, [0] * len(ready), [2 ** 62])) == {0}
    # mixed dtypes never share a bucket
    assert plan_buckets([8, 8, 8, 8], [0, 1, 0, 1], [2 ** 62]) == [0, 1, 0, 1]


def test_state_dict_keys_unwrapped():
    sd = ConvNet().state_dict()
    assert {k: tuple(v.shape) for k, v in sd.items()} == REF_KEYS
    assert sd["batchnorm.num_batches_tracked"].dtype == torch.int64
    # loads into the reference architecture built from stock torch layers
    class RefNet(nn.Module):
        def __init__(self):
            super().__init__()
            self.conv1 = nn.Conv2d(1, 32, 3, 1); self.conv2 = nn.Conv2d(32, 64, 3, 1)
            self.dropout1 = nn.Dropout2d(0.25); self.dropout2 = nn.Dropout2d(0.5)
            self.fc1 = nn.Linear(9216, 128); self.fc2 = nn.Linear(128, 10); self.batchnorm = nn.BatchNorm1d(128)
    RefNet().load_state_dict(sd)


class TinyNet(nn.Module):
    """No BatchNorm, so N ranks x batch B must equal 1 rank x batch N*B exactly (up to fp rounding)."""
    def __init__(self):
        super().__init__()
        self.conv = nn.Conv2d(1, 4, 3); self.fc1 = nn.Linear(4 * 6 * 6, 300); self.fc2 = nn.Linear(300, 10)
        self.unused = None
    def forward(self, x):
        x = F.relu(self.conv(x)); x = F.max_pool2d(x, 2).flatten(1)
        return self.fc2(F.relu(self.fc1(x)))


def _train_worker(rank, world, port, tmp):
    dist = init(rank, world, port, "gloo")
    torch.manual_seed(0)
    model = TinyNet()
    ref = TinyNet(); ref.load_state_dict(model.state_dict())
    if rank == 1:  # DDP must overwrite this with rank 0's values
        with torch.no_grad():
            for p in model.parameters(): p.add_(1.0)
    ddp = dcp.DistributedDataParallel(model, bucket_cap_mb=0.002)
    g = torch.Generator().manual_seed(7)
    X, Y = torch.randn(4, 8, 1, 14, 14, generator=g), torch.randint(0, 10, (4, 8), generator=g)
    opt = dcp.ops.Adadelta(ddp.parameters(), lr=0.5)
    ropt = torch.optim.Adadelta(ref.parameters(), lr=0.5)
    half = 8 // world
    for step in range(4):
        opt.zero_grad()
        loss = F.cross_entropy(ddp(X[step, rank * half:(rank + 1) * half]), Y[step, rank * half:(rank + 1) * half])
        loss.backward()
        ropt.zero_grad()
        F.cross_entropy(ref(X[step]), Y[step]).backward()
        for (n, p), q in zip(model.named_parameters(), ref.parameters()):
            assert torch.allclose(p.grad, q.grad, atol=1e-6, rtol=1e-4), (step, n)
        opt.step(); ropt.step()
        for (n, p), q in zip(model.named_parameters(), ref.parameters()):
            assert torch.allclose(p, q, atol=1e-5, rtol=1e-4), (step, n)
        if step == 0:
            assert len(ddp.bucket_sizes_bytes()) == 1
    assert len(ddp.bucket_sizes_bytes()) > 1  # rebuilt after iteration 1 with the small cap
    # replicas are bit-identical
    flat = torch.cat([p.detach().reshape(-1) for p in model.parameters()])
    both = [torch.empty_like(flat) for _ in range(world)]
    dist.all_gather(both, flat)
    assert all(torch.equal(both[0], b) for b in both)
    # no_sync accumulates locally, the next synced backward reduces the sum
    opt.zero_grad()
    with ddp.no_sync():
        F.cross_entropy(ddp(X[0, rank * half:(rank + 1) * half]), Y[0, rank * half:(rank + 1) * half]).backward()
    local = [p.grad.clone() for p in model.parameters()]
    gathered = [torch.empty_like(local[0]) for _ in range(world)]
    dist.all_gather(gathered, local[0])
    assert not torch.equal(gathered[0], gathered[1])
    F.cross_entropy(ddp(X[1, rank * half:(rank + 1) * half]), Y[1, rank * half:(rank + 1) * half]).backward()
    ropt.zero_grad()
    (F.cross_entropy(ref(X[0]), Y[0]) + F.cross_entropy(ref(X[1]), Y[1])).backward()
    for p, q in zip(model.parameters(), ref.parameters()):
        assert torch.allclose(p.grad, q.grad, atol=1e-5, rtol=1e-4)
    sd = ddp.state_dict()
    assert all(k.startswith("module.") for k in sd)
    if rank == 0:
        torch.save(sd, os.path.join(tmp, "ck.pt"))
    dist.barrier()
    loaded = torch.load(os.path.join(tmp, "ck.pt"))
    assert list(loaded.keys()) == list(sd.keys())
    dist.destroy_process_group()


def test_ddp_matches_single_process_large_batch():
    with tempfile.TemporaryDirectory() as tmp:
        dcp.spawn(_train_worker, args=(2, free_port(), tmp), nprocs=2, join=True)


def _unused_worker(rank, world, port):
    dist = init(rank, world, port, "gloo")
    class Net(nn.Module):
        def __init__(self):
            super().__init__(); self.a = nn.Linear(4, 4); self.b = nn.Linear(4, 4)
        def forward(self, x): return self.a(x)
    ddp = dcp.DistributedDataParallel(Net())
    with pytest.raises(RuntimeError, match="find_unused_parameters"):
        ddp(torch.randn(2, 4)).sum().backward()
    ddp2 = dcp.DistributedDataParallel(Net(), find_unused_parameters=True)
    ddp2(torch.randn(2, 4)).sum().backward()
    assert ddp2.module.a.weight.grad is not None
    dist.destroy_process_group()


def test_unused_parameters():
    dcp.spawn(_unused_worker, args=(2, free_port()), nprocs=2, join=True)


def _buffers_worker(rank, world, port):
    dist = init(rank, world, port, "gloo")
    torch.manual_seed(rank)  # different data per rank -> different local BN stats
    ddp = dcp.DistributedDataParallel(ConvNet())
    ddp.train()
    out = ddp(torch.randn(4, 1, 28, 28))
    dcp.ops.nll_loss(out, torch.randint(0, 10, (4,))).backward()
    ddp(torch.randn(4, 1, 28, 28))  # forward #2 re-broadcasts rank 0's running stats first (C4)
    ddp.eval()
    with torch.no_grad():
        ddp(torch.randn(4, 1, 28, 28))  # first eval forward syncs the buffers once more
    rm = ddp.module.batchnorm.running_mean.clone()
    both = [torch.empty_like(rm) for _ in range(world)]
    dist.all_gather(both, rm)
    assert torch.equal(both[0], both[1])
    dist.destroy_process_group()


def test_buffers_follow_rank0():
    dcp.spawn(_buffers_worker, args=(2, free_port()), nprocs=2, join=True)
