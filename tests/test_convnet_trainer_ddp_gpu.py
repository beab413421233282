"""FusedConvNetTrainer under DistributedDataParallel on >= 2 GPUs: after a few captured steps on different data per
rank the replicas hold identical parameters (bucketed all-reduce of the gradients, split optimizer update), the
BatchNorm buffers follow rank 0 (per-forward buffer broadcast, torch DDP semantics), and the averaged gradient equals
the mean of the per-rank gradients computed without DDP."""
import pytest
import torch

import distributed_compute_pytorch_b200 as dcp
from tests.helpers import free_port, init

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]


def _worker(rank, world, port):
    from distributed_compute_pytorch_b200 import ops
    from distributed_compute_pytorch_b200.parallel.convnet_trainer import FusedConvNetTrainer
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist = init(rank, world, port, "nccl")
    N = 64
    torch.manual_seed(0)
    g = torch.Generator().manual_seed(100 + rank)
    xs = torch.randn(6, N, 1, 28, 28, generator=g).to(dev); ys = torch.randint(0, 10, (6, N), generator=g).to(dev)

    # per-rank gradients without DDP (dropout off), to check the averaging
    torch.manual_seed(0)
    solo = dcp.models.ConvNet().to(dev).train(); solo.dropout1.p = solo.dropout2.p = 0.0
    tr0 = FusedConvNetTrainer(solo, ops.Adadelta(solo.parameters(), lr=0.0), N, use_graph=False)
    tr0.step(xs[0], ys[0]); torch.cuda.synchronize()
    local = torch.cat([p.grad.reshape(-1) for p in solo.parameters()]).clone()
    mean_grad = local.clone(); dist.all_reduce(mean_grad); mean_grad /= world

    torch.manual_seed(0)
    net = dcp.models.ConvNet().to(dev).train(); net.dropout1.p = net.dropout2.p = 0.0
    ddp = dcp.DistributedDataParallel(net, device_ids=[rank])
    opt = ops.Adadelta(ddp.parameters(), lr=0.0)
    tr = FusedConvNetTrainer(ddp, opt, N, use_graph=False)
    tr.step(xs[0], ys[0]); torch.cuda.synchronize()
    got = torch.cat([p.grad.reshape(-1) for p in net.parameters()])
    err = ((got - mean_grad).norm() / mean_grad.norm()).item()
    assert err < 1e-5, err
    # second step on the same batch (lr = 0: same weights): the optimizer's pointer table exists now, so the last bucket goes
    # through the fused tail - producers push into every rank's payload block, one kernel waits, averages and updates
    assert tr.dp_tail and tr._opt_rows() is not None
    for q in net.parameters():
        q.grad.zero_()
    tr.step(xs[0], ys[0]); torch.cuda.synchronize()
    got = torch.cat([p.grad.reshape(-1) for p in net.parameters()])
    err = ((got - mean_grad).norm() / mean_grad.norm()).item()
    assert err < 1e-5, err
    ref_g = got.clone(); dist.broadcast(ref_g, 0)
    assert torch.equal(got, ref_g), "averaged gradients differ between ranks"

    # captured steps with a real learning rate: replicas stay identical, buffers follow rank 0
    for grp in opt.param_groups: grp["lr"] = 1.0
    tr._capture()
    if rank != 0:                                   # perturb a non-root replica's running stats: the broadcast must repair them
        net.batchnorm.running_mean.add_(5.0)
    for i in range(1, 6):
        tr.step(xs[i], ys[i])
    torch.cuda.synchronize()
    flat = torch.cat([p.detach().reshape(-1) for p in net.parameters()])
    ref = flat.clone(); dist.broadcast(ref, 0)
    assert torch.equal(flat, ref), (flat - ref).abs().max().item()
    rm = net.batchnorm.running_mean.clone(); rm0 = rm.clone(); dist.broadcast(rm0, 0)
    # rank r's buffer = momentum-blend of rank 0's previous buffer with rank r's own last batch statistics
    # rank 0's head kernel writes its running statistics into every rank's buffer block: replicas hold rank 0's values
    assert torch.equal(rm, rm0), (rm - rm0).abs().max().item()
    nbt = net.batchnorm.num_batches_tracked.clone(); nbt0 = nbt.clone(); dist.broadcast(nbt0, 0)
    assert nbt.item() == nbt0.item()
    assert torch.isfinite(flat).all()

    # the library reducer path (autograd hooks) with per-bucket collective timing switched on
    torch.manual_seed(0)
    net2 = dcp.models.ConvNet().to(dev).train()
    ddp2 = dcp.DistributedDataParallel(net2, device_ids=[rank])
    ddp2.enable_comm_timing()
    for i in range(2):
        ddp2.zero_grad()
        ops.nll_loss(ddp2(xs[i]), ys[i]).backward()
    rec = ddp2.comm_timings()
    assert len(rec) >= 3 and all(ms >= 0.0 and nbytes > 0 for _, nbytes, ms in rec), rec       # 1 bucket in step 1, 2 after the rebuild
    assert max(nb for _, nb, _ in rec) >= 4725288        # the big bucket (fc1.weight & co.) went through the timed path
    dist.barrier()
    dist.destroy_process_group()


def test_fused_trainer_under_ddp():
    world = min(torch.cuda.device_count(), 4)
    dcp.spawn(_worker, args=(world, free_port()), nprocs=world, join=True)
