"""SyncBatchNorm: statistics reduced across ranks with the one-shot P2P all_reduce.  W ranks x batch B must equal a
single-process BatchNorm over the concatenated batch (forward, input gradient, parameter gradients after DDP averaging)."""
import pytest
import torch

import distributed_compute_pytorch_b200 as dcp
from distributed_compute_pytorch_b200 import ops
from tests.helpers import free_port, init

pytestmark = [pytest.mark.gpu]


def _worker(rank, world, port):
    torch.cuda.set_device(rank)
    dist = init(rank, world, port, "nccl")
    dev = torch.device("cuda", rank)
    g = torch.Generator().manual_seed(3)
    for shape, relu in (((8 * world, 64, 6, 6), True), ((16 * world, 128), False)):
        X = (torch.randn(*shape, generator=g) * 2 + 0.5)
        G = torch.randn(*shape, generator=g)
        C = shape[1]
        ref = (ops.BatchNorm2d if len(shape) == 4 else ops.BatchNorm1d)(C, fuse_relu=relu).to(dev)
        with torch.no_grad():   # same affine parameters on every rank (as DDP guarantees): draw them from the shared CPU generator
            ref.weight.copy_(torch.rand(C, generator=g) + 0.5); ref.bias.copy_(torch.rand(C, generator=g) - 0.5)
        sbn = ops.SyncBatchNorm(C, fuse_relu=relu).to(dev); sbn.load_state_dict(ref.state_dict())
        xr = X.to(dev).requires_grad_(True)
        if len(shape) == 4:
            xr = xr.detach().contiguous(memory_format=torch.channels_last).requires_grad_(True)
        yr = ref(xr); yr.backward(G.to(dev))
        n = shape[0] // world
        xl = X[rank * n:(rank + 1) * n].to(dev)
        if len(shape) == 4:
            xl = xl.contiguous(memory_format=torch.channels_last)
        xl.requires_grad_(True)
        yl = sbn(xl); yl.backward(G[rank * n:(rank + 1) * n].to(dev))
        assert torch.allclose(yl, yr[rank * n:(rank + 1) * n], rtol=1e-4, atol=1e-4)
        assert torch.allclose(xl.grad, xr.grad[rank * n:(rank + 1) * n], rtol=1e-3, atol=1e-4)
        # parameter gradients are local sums; their SUM over ranks equals the full-batch gradient
        for a, b in ((sbn.weight.grad, ref.weight.grad), (sbn.bias.grad, ref.bias.grad)):
            t = a.clone(); dist.all_reduce(t)
            assert torch.allclose(t, b, rtol=1e-3, atol=1e-3)
        assert torch.allclose(sbn.running_mean, ref.running_mean, rtol=1e-4, atol=1e-5)
        assert torch.allclose(sbn.running_var, ref.running_var, rtol=1e-3, atol=1e-4)
    m = ops.SyncBatchNorm.convert_sync_batchnorm(dcp.models.resnet18(num_classes=10))
    assert isinstance(m.bn1, ops.SyncBatchNorm) and isinstance(m.layer1[0].bn2, ops.SyncBatchNorm) and m.bn1.fuse_relu
    dist.destroy_process_group()


@pytest.mark.multigpu
def test_sync_batchnorm_equals_full_batch():
    world = min(torch.cuda.device_count(), 4)
    dcp.spawn(_worker, args=(world, free_port()), nprocs=world, join=True)
