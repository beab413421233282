"""Every primitive over the CPU mesh, world_size 2 and 3, checked against closed-form expectations
(the same values a gloo run of the reference's backend produces)."""
import pytest
import torch

import distributed_compute_pytorch_b200 as dcp
from tests.helpers import free_port, init


def _worker(rank, world, port):
    dist = init(rank, world, port, "gloo")
    assert dist.get_rank() == rank and dist.get_world_size() == world and dist.is_initialized()
    R = dist.ReduceOp
    # all_reduce: dtypes, sizes (small path + ring path), odd sizes, 0-dim
    for dtype in (torch.float32, torch.float64, torch.int64, torch.int32, torch.bfloat16):
        for n in (1, 7, 1000, 100003):
            t = (torch.arange(n) % 13 + rank).to(dtype)
            dist.all_reduce(t, op=R.SUM)
            exp = sum(((torch.arange(n) % 13 + r).to(dtype) for r in range(world)))
            assert torch.equal(t, exp), (dtype, n)
    z = torch.tensor(float(rank + 1), requires_grad=True) * 2.0  # 0-dim non-leaf needing grad (ref:main.py:65)
    dist.all_reduce(z, op=R.SUM)
    assert z.item() == 2.0 * sum(range(1, world + 1))
    c = torch.tensor(rank + 5, dtype=torch.int64)  # 0-dim int64 (ref:main.py:91)
    dist.all_reduce(c)
    assert c.item() == sum(r + 5 for r in range(world))
    for op, fn in ((R.MAX, max), (R.MIN, min)):
        t = torch.tensor([float(rank), float(-rank)])
        dist.all_reduce(t, op=op)
        assert t.tolist() == [fn(float(r) for r in range(world)), fn(float(-r) for r in range(world))]
    t = torch.full((50000,), float(rank + 1)); dist.all_reduce(t, op=R.AVG)
    assert torch.allclose(t, torch.full((50000,), (world + 1) / 2))
    t = torch.tensor([rank + 2.0]); dist.all_reduce(t, op=R.PRODUCT)
    exp = 1.0
    for r in range(world): exp *= r + 2.0
    assert t.item() == exp
    # non-contiguous
    m = torch.ones(4, 6)[:, ::2] * (rank + 1)
    dist.all_reduce(m)
    assert torch.equal(m, torch.full((4, 3), float(sum(range(1, world + 1)))))
    # broadcast from every root
    for src in range(world):
        t = torch.arange(33, dtype=torch.float32) * (rank + 1)
        dist.broadcast(t, src=src)
        assert torch.equal(t, torch.arange(33, dtype=torch.float32) * (src + 1))
    # reduce to a non-zero root
    dst = world - 1
    t = torch.full((9,), float(rank + 1)); dist.reduce(t, dst=dst)
    if rank == dst:
        assert torch.equal(t, torch.full((9,), float(sum(range(1, world + 1)))))
    # all_gather / all_gather_into_tensor
    outs = [torch.empty(5) for _ in range(world)]
    dist.all_gather(outs, torch.full((5,), float(rank)))
    for r in range(world): assert torch.equal(outs[r], torch.full((5,), float(r)))
    flat = torch.empty(world * 3); dist.all_gather_into_tensor(flat, torch.full((3,), float(rank)))
    assert torch.equal(flat, torch.arange(world).repeat_interleave(3).float())
    # gather / scatter
    gl = [torch.empty(2) for _ in range(world)] if rank == 1 else None
    dist.gather(torch.tensor([rank, rank * 10.0]), gl, dst=1)
    if rank == 1:
        for r in range(world): assert gl[r].tolist() == [r, r * 10.0]
    sl = [torch.full((4,), float(r * 3)) for r in range(world)] if rank == 0 else None
    t = torch.empty(4); dist.scatter(t, sl, src=0)
    assert torch.equal(t, torch.full((4,), float(rank * 3)))
    # reduce_scatter / reduce_scatter_tensor
    out = torch.empty(3)
    dist.reduce_scatter(out, [torch.full((3,), float(rank + i)) for i in range(world)])
    assert torch.equal(out, torch.full((3,), float(sum(r + rank for r in range(world)))))
    out = torch.empty(2); dist.reduce_scatter_tensor(out, torch.arange(2 * world, dtype=torch.float32) + rank)
    exp = sum(torch.arange(2 * world, dtype=torch.float32) + r for r in range(world))[2 * rank: 2 * rank + 2]
    assert torch.equal(out, exp)
    # send / recv ring
    nxt, prv = (rank + 1) % world, (rank - 1) % world
    buf = torch.empty(1000)
    if rank % 2 == 0:
        dist.send(torch.full((1000,), float(rank)), dst=nxt); dist.recv(buf, src=prv)
    else:
        dist.recv(buf, src=prv); dist.send(torch.full((1000,), float(rank)), dst=nxt)
    assert torch.equal(buf, torch.full((1000,), float(prv)))
    w = dist.all_reduce(torch.ones(3), async_op=True); w.wait(); assert w.is_completed()
    dist.barrier()
    dist.destroy_process_group()
    assert not dist.is_initialized()


@pytest.mark.parametrize("world", [2, 3])
def test_all_primitives(world):
    dcp.spawn(_worker, args=(world, free_port()), nprocs=world, join=True)


def _exchange(rank, world, port):
    """The exchange every rank runs at once - post irecv, post isend, wait both - needs receives and sends to progress
    independently; with blocking stand-ins it deadlocks.  8 MB messages do not fit a socket buffer."""
    dist = init(rank, world, port, "gloo")
    nxt, prv = (rank + 1) % world, (rank - 1) % world
    for n in (5, 2_000_000):
        inbox = torch.empty(n)
        out = torch.full((n,), float(rank))
        r = dist.irecv(inbox, src=prv)
        s = dist.isend(out, dst=nxt)
        assert r.wait() and s.wait()
        assert r.is_completed() and s.is_completed()
        assert torch.equal(inbox, torch.full((n,), float(prv)))
    # several operations in flight, completed out of posting order, followed by a collective (which drains whatever is pending)
    bufs = [torch.empty(1000) for _ in range(3)]
    rs = [dist.irecv(b, src=prv, tag=k) for k, b in enumerate(bufs)]
    ss = [dist.isend(torch.full((1000,), float(10 * rank + k)), dst=nxt, tag=k) for k in range(3)]
    t = torch.ones(4); dist.all_reduce(t)
    assert t[0].item() == world
    for w in reversed(rs + ss):
        w.wait()
    for k, b in enumerate(bufs):
        assert torch.equal(b, torch.full((1000,), float(10 * prv + k)))
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_isend_irecv_exchange_does_not_deadlock(world):
    dcp.spawn(_exchange, args=(world, free_port()), nprocs=world, join=True)


def _failing(rank, world, port):
    dist = init(rank, world, port, "gloo")
    if rank == 1:
        raise ValueError("boom from rank 1")
    import time
    time.sleep(60)


def test_spawn_propagates_first_failure_and_kills_siblings():
    import time
    t0 = time.time()
    with pytest.raises(dcp.ProcessRaisedException) as ei:
        dcp.spawn(_failing, args=(2, free_port()), nprocs=2, join=True)
    assert "boom from rank 1" in str(ei.value) and ei.value.error_index == 1
    assert time.time() - t0 < 45


def test_errors_without_init():
    with pytest.raises(RuntimeError):
        dcp.all_reduce(torch.ones(1))
    with pytest.raises(RuntimeError):
        dcp.get_rank()
