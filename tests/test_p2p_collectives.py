"""Hand-written NVLink collectives vs closed-form expectations (the values a gloo-CPU run gives),
every primitive x dtype x size x algorithm, plus repeated-call stress for the epoch/parity protocol."""
import os

import pytest
import torch

import distributed_compute_pytorch_b200 as dcp
from tests.helpers import free_port, init

pytestmark = [pytest.mark.gpu]


def _ngpu():
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


def _pattern(n, rank, dtype, device):
    base = (torch.arange(n, device=device) % 251).to(torch.float32) * 0.25 + rank
    return base.to(dtype)


def _worker(rank, world, port, symm_mode):
    os.environ["DCP_SYMM"] = symm_mode
    os.environ["DCP_COLL_TIMEOUT_S"] = "20"
    torch.cuda.set_device(rank)
    dist = init(rank, world, port, "nccl")
    dev = torch.device("cuda", rank)
    be = dist.distributed._world().cuda
    R = dist.ReduceOp
    algos = [1, 2] + ([3] if be.nvls else [])
    sizes = [1, 3, 257, 4096, 65536 + 5, 1 << 20, (1 << 22) + 3]
    for dtype in (torch.float32, torch.bfloat16, torch.float16, torch.int64, torch.int32, torch.float64):
        for n in sizes:
            exp = sum(_pattern(n, r, torch.float64 if dtype.is_floating_point else dtype, dev) for r in range(world))
            for algo in algos:
                if algo == 3 and dtype not in (torch.float32, torch.bfloat16, torch.float16):
                    continue
                if algo == 1 and n * torch.empty((), dtype=dtype).element_size() > be.oneshot_max:
                    continue
                for symmetric in (False, True):
                    if symmetric and algo == 1:
                        continue
                    t = be.empty_symmetric(n, dtype) if symmetric else torch.empty(n, dtype=dtype, device=dev)
                    for rep in range(3):  # repeated calls exercise parity / epoch reuse
                        t.copy_(_pattern(n, rank, dtype, dev))
                        be.all_reduce(t, R.SUM, algo=algo)
                    tol = dict(rtol=2e-2, atol=2e-2) if dtype in (torch.bfloat16, torch.float16) else dict(rtol=1e-6, atol=1e-6)
                    if dtype.is_floating_point:
                        assert torch.allclose(t.double(), exp.double(), **tol), (dtype, n, algo, symmetric)
                    else:
                        assert torch.equal(t, exp), (dtype, n, algo, symmetric)
                    if symmetric:
                        be.free_symmetric(t)
    # default dispatcher, AVG / MAX / MIN / PRODUCT, 0-dim tensors (ref:main.py:65,90,91)
    t = torch.full((100000,), float(rank + 1), device=dev); dist.all_reduce(t, op=R.AVG)
    assert torch.allclose(t, torch.full_like(t, (world + 1) / 2))
    t = torch.tensor([float(rank), float(-rank)], device=dev); dist.all_reduce(t, op=R.MAX)
    assert t.tolist() == [float(world - 1), 0.0]
    t = torch.tensor([float(rank), float(-rank)], device=dev); dist.all_reduce(t, op=R.MIN)
    assert t.tolist() == [0.0, float(-(world - 1))]
    z = (torch.tensor(float(rank + 1), device=dev, requires_grad=True) * 2.0)
    dist.all_reduce(z, op=R.SUM); assert z.item() == 2.0 * sum(range(1, world + 1))
    c = torch.tensor(rank + 5, dtype=torch.int64, device=dev); dist.all_reduce(c)
    assert c.item() == sum(r + 5 for r in range(world))
    m = torch.ones(8, 6, device=dev)[:, ::2] * (rank + 1); dist.all_reduce(m)
    assert torch.equal(m, torch.full((8, 3), float(sum(range(1, world + 1))), device=dev))
    # replicas bit-identical after a float all_reduce
    t = torch.randn(300001, device=dev, generator=torch.Generator(device=dev).manual_seed(rank)); dist.all_reduce(t)
    both = [torch.empty_like(t) for _ in range(world)]; dist.all_gather(both, t)
    assert all(torch.equal(both[0], b) for b in both)
    # broadcast from every root, staged and symmetric
    for src in range(world):
        for n in (1, 1000, (1 << 20) + 7):
            t = _pattern(n, rank, torch.float32, dev); dist.broadcast(t, src=src)
            assert torch.equal(t, _pattern(n, src, torch.float32, dev))
        s = be.empty_symmetric(5000, torch.float32); s.copy_(_pattern(5000, rank, torch.float32, dev)); dist.broadcast(s, src=src)
        assert torch.equal(s, _pattern(5000, src, torch.float32, dev)); be.free_symmetric(s)
    # reduce to non-zero root
    dst = world - 1
    t = torch.full((70001,), float(rank + 1), device=dev); dist.reduce(t, dst=dst)
    if rank == dst:
        assert torch.equal(t, torch.full_like(t, float(sum(range(1, world + 1)))))
    # all_gather / into_tensor / gather / scatter / reduce_scatter
    for n in (5, 100003):
        outs = [torch.empty(n, device=dev) for _ in range(world)]
        dist.all_gather(outs, torch.full((n,), float(rank), device=dev))
        for r in range(world): assert torch.equal(outs[r], torch.full((n,), float(r), device=dev))
    so = be.empty_symmetric(world * 4096, torch.float32); dist.all_gather_into_tensor(so, torch.full((4096,), float(rank), device=dev))
    assert torch.equal(so, torch.arange(world, device=dev).repeat_interleave(4096).float()); be.free_symmetric(so)
    gl = [torch.empty(33, device=dev) for _ in range(world)] if rank == 1 else None
    dist.gather(torch.full((33,), float(rank * 10), device=dev), gl, dst=1)
    if rank == 1:
        for r in range(world): assert torch.equal(gl[r], torch.full((33,), float(r * 10), device=dev))
    sl = [torch.full((77,), float(r * 3), device=dev) for r in range(world)] if rank == 0 else None
    t = torch.empty(77, device=dev); dist.scatter(t, sl, src=0)
    assert torch.equal(t, torch.full((77,), float(rank * 3), device=dev))
    for n in (3, 50001):
        out = torch.empty(n, device=dev)
        dist.reduce_scatter(out, [torch.full((n,), float(rank + i), device=dev) for i in range(world)])
        assert torch.equal(out, torch.full((n,), float(sum(r + rank for r in range(world))), device=dev))
    out = torch.empty(1024, device=dev, dtype=torch.bfloat16)
    dist.reduce_scatter_tensor(out, torch.ones(1024 * world, device=dev, dtype=torch.bfloat16) * (rank + 1), op=R.AVG)
    assert torch.allclose(out.float(), torch.full((1024,), (world + 1) / 2, device=dev))
    # send / recv ring, small and multi-slot messages
    nxt, prv = (rank + 1) % world, (rank - 1) % world
    for n in (1, 1000, 3 * (1 << 20) + 1):
        buf = torch.empty(n, device=dev)
        if rank % 2 == 0:
            dist.send(torch.full((n,), float(rank), device=dev), dst=nxt); dist.recv(buf, src=prv)
        else:
            dist.recv(buf, src=prv); dist.send(torch.full((n,), float(rank), device=dev), dst=nxt)
        assert torch.equal(buf, torch.full((n,), float(prv), device=dev))
    # the same ring with EVERY rank posting irecv, then isend, then waiting (receives and sends run on separate side streams;
    # with blocking stand-ins each rank's receive kernel would sit in front of its own send)
    for n in (1000, 3 * (1 << 20) + 1):
        buf = torch.empty(n, device=dev); out = torch.full((n,), float(rank), device=dev)
        r = dist.irecv(buf, src=prv); s = dist.isend(out, dst=nxt)
        r.wait(); s.wait()
        assert torch.equal(buf, torch.full((n,), float(prv), device=dev))
        torch.cuda.synchronize(); assert r.is_completed() and s.is_completed()
    # ... and a blocking pair right behind the non-blocking ones (ordered after them)
    buf = torch.empty(77, device=dev)
    if rank % 2 == 0:
        dist.send(torch.full((77,), float(rank), device=dev), dst=nxt); dist.recv(buf, src=prv)
    else:
        dist.recv(buf, src=prv); dist.send(torch.full((77,), float(rank), device=dev), dst=nxt)
    assert torch.equal(buf, torch.full((77,), float(prv), device=dev))
    w = dist.all_reduce(torch.ones(3, device=dev), async_op=True); w.wait(); torch.cuda.synchronize(); assert w.is_completed()
    # stress: 200 back-to-back small all_reduces without host sync
    acc = torch.zeros(64, device=dev)
    for i in range(200):
        t = torch.full((64,), float(i % 7), device=dev); dist.all_reduce(t); acc += t
    assert torch.equal(acc, torch.full((64,), float(world * sum(i % 7 for i in range(200))), device=dev))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.multigpu
@pytest.mark.parametrize("symm_mode", ["vmm", "ipc"])
def test_all_primitives_multi_gpu(symm_mode):
    world = min(_ngpu(), 8)
    dcp.spawn(_worker, args=(world, free_port(), symm_mode), nprocs=world, join=True)


def _single(rank, world, port):
    torch.cuda.set_device(0)
    dist = init(rank, world, port, "nccl")
    be = dist.distributed._world().cuda
    t = be.empty_symmetric(1000, torch.float32); t.fill_(3.0)
    assert be.is_symmetric(t) and not be.is_symmetric(torch.zeros(1, device="cuda"))
    dist.all_reduce(t); assert torch.equal(t, torch.full_like(t, 3.0))
    x = torch.arange(10, device="cuda", dtype=torch.float32); dist.broadcast(x, 0); dist.barrier()
    be.free_symmetric(t)
    t2 = be.empty_symmetric(1000, torch.float32)
    assert t2.data_ptr() == t.data_ptr()  # allocator reuses the freed block
    dist.destroy_process_group()


def test_world_size_one_backend():
    dcp.spawn(_single, args=(1, free_port()), nprocs=1, join=True)


def _channels_worker(rank, world, port):
    """One-shot all_reduces of ordinary tensors on three channels from three streams AT ONCE (what a small DDP bucket on the
    communication stream, a SyncBatchNorm statistic on the compute stream and a user collective amount to): every channel owns
    its staging slots, so nobody's payload lands in somebody else's."""
    from distributed_compute_pytorch_b200.parallel.p2p_backend import CH_BN, CH_DDP, CH_USER
    torch.cuda.set_device(rank)
    dist = init(rank, world, port, "nccl")
    dev = torch.device("cuda", rank)
    be = dist.distributed._world().cuda
    R = dist.ReduceOp
    streams = [torch.cuda.Stream(dev) for _ in range(3)]
    chans = (CH_DDP, CH_BN, CH_USER)
    n = 2048
    for it in range(100):
        ts = []
        for k, (st, ch) in enumerate(zip(streams, chans)):
            with torch.cuda.stream(st):
                t = torch.full((n,), float((rank + 1) * (k + 1) + it), device=dev)
                be.all_reduce(t, R.SUM, stream=st, algo=1, chan=ch)
                ts.append(t)
        torch.cuda.synchronize()
        for k, t in enumerate(ts):
            exp = float(sum((r + 1) * (k + 1) + it for r in range(world)))
            assert torch.equal(t, torch.full((n,), exp, device=dev)), (it, k, t[:4].tolist(), exp)
    be.check_health()
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.multigpu
def test_concurrent_channels_do_not_share_staging():
    world = min(_ngpu(), 4)
    dcp.spawn(_channels_worker, args=(world, free_port()), nprocs=world, join=True)
