"""The multi-rank collective kernels on ONE GPU: every rank of a small world runs inside one cooperative launch
(parallel/emulated.py, csrc/coll/coll_launch.cuh), so the signal protocol, the parity double-buffering, the chunked staged
paths and the reductions of csrc/coll are exercised by the single-GPU test tier.  Expected values are closed forms of the
per-rank patterns (what a gloo-CPU run of the same collective gives); the multi-GPU tests (test_p2p_collectives.py) run the
same checks over real NVLink, NVLS paths included.
"""
import ctypes as C

import pytest
import torch

from distributed_compute_pytorch_b200.parallel.emulated import EmulatedWorld
from distributed_compute_pytorch_b200.parallel.p2p_backend import CH_OPT, CH_USER
from distributed_compute_pytorch_b200.types import ReduceOp as R

pytestmark = pytest.mark.gpu
CTAS = 4        # per rank; world x CTAS blocks must be co-resident


def _pattern(n, rank, dtype, dev):
    base = (torch.arange(n, device=dev) % 251).to(torch.float32) * 0.25 + rank
    return base.to(dtype)


def _tol(dtype):
    return dict(rtol=2e-2, atol=2e-2) if dtype in (torch.bfloat16, torch.float16) else dict(rtol=1e-6, atol=1e-6)


@pytest.mark.parametrize("world", [2, 4, 8])
def test_all_reduce_every_algorithm_dtype_and_size(world):
    dev = torch.device("cuda", 0)
    with EmulatedWorld(world, staging_mb=32) as w:
        for dtype in (torch.float32, torch.bfloat16, torch.int64, torch.float64):
            es = torch.empty((), dtype=dtype).element_size()
            for n in (1, 3, 257, 65536 + 5, (1 << 20) + 3):
                exp = sum(_pattern(n, r, torch.float64 if dtype.is_floating_point else dtype, dev) for r in range(world))
                for algo in (1, 2):
                    if algo == 1 and n * es > w.ranks[0].oneshot_max:
                        continue
                    for symmetric in (False, True):
                        if symmetric and algo == 1:
                            continue
                        ts = [be.empty_symmetric(n, dtype) if symmetric else torch.empty(n, dtype=dtype, device=dev) for be in w.ranks]
                        for rep in range(3):            # repeated calls: launch-counter parity, epoch reuse
                            for r, be in enumerate(w.ranks):
                                ts[r].copy_(_pattern(n, r, dtype, dev))
                                be.all_reduce(ts[r], R.SUM, algo=algo, max_ctas=CTAS)
                            w.flush()
                        torch.cuda.synchronize(); w.check_health()
                        for t in ts:
                            if dtype.is_floating_point:
                                assert torch.allclose(t.double(), exp.double(), **_tol(dtype)), (dtype, n, algo, symmetric)
                            else:
                                assert torch.equal(t, exp), (dtype, n, algo, symmetric)
                        if symmetric:
                            for be, t in zip(w.ranks, ts):
                                be.free_symmetric(t)
        # AVG / MAX on the default dispatcher, 0-dim tensors (the reference's calls, ref:main.py:65,90,91)
        ts = [torch.full((1000,), float(r + 1), device=dev) for r in range(world)]
        for be, t in zip(w.ranks, ts):
            be.all_reduce(t, R.AVG, max_ctas=CTAS)
        cs = [torch.tensor(r + 5, dtype=torch.int64, device=dev) for r in range(world)]
        for be, t in zip(w.ranks, cs):
            be.all_reduce(t, R.SUM, max_ctas=CTAS)
        ms = [torch.tensor([float(r), float(-r)], device=dev) for r in range(world)]
        for be, t in zip(w.ranks, ms):
            be.all_reduce(t, R.MAX, max_ctas=CTAS)
        w.flush(); torch.cuda.synchronize(); w.check_health()
        assert all(torch.allclose(t, torch.full_like(t, (world + 1) / 2)) for t in ts)
        assert all(t.item() == sum(r + 5 for r in range(world)) for t in cs)
        assert all(t.tolist() == [float(world - 1), 0.0] for t in ms)


@pytest.mark.parametrize("world", [2, 4])
def test_rooted_and_gather_scatter_collectives_with_chunking(world):
    """Small staging area (8 MiB) so that multi-megabyte shards go through the chunked loops of the staged paths."""
    dev = torch.device("cuda", 0)
    with EmulatedWorld(world, staging_mb=8) as w:
        lib = w.lib
        st = torch.cuda.current_stream().cuda_stream
        cap = lib.dcp_coll_bounce_max_bytes(w.ranks[0].handle)
        big = (cap // 4) * 3 // world + 1237            # elements: > one bounce chunk per shard for all_gather / reduce_scatter
        for n in (1, 1000, big):
            # broadcast from every root, staged and symmetric
            for root in range(world):
                ts = [_pattern(n, r, torch.float32, dev) for r in range(world)]
                for be, t in zip(w.ranks, ts):
                    be.broadcast(t, root)
                w.flush(); torch.cuda.synchronize()
                assert all(torch.equal(t, _pattern(n, root, torch.float32, dev)) for t in ts), ("broadcast", n, root)
            # all_gather_into_tensor
            ins = [_pattern(n, r, torch.float32, dev) for r in range(world)]
            outs = [torch.full((world * n,), -1.0, device=dev) for _ in range(world)]
            for be, o, i_ in zip(w.ranks, outs, ins):
                be.all_gather_into_tensor(o, i_)
            w.flush(); torch.cuda.synchronize()
            exp = torch.cat(ins)
            assert all(torch.equal(o, exp) for o in outs), ("all_gather", n)
            # reduce_scatter_tensor (SUM) and reduce to a non-zero root
            full = [(_pattern(world * n, r, torch.float32, dev)) for r in range(world)]
            outs = [torch.empty(n, device=dev) for _ in range(world)]
            for be, o, i_ in zip(w.ranks, outs, full):
                be.reduce_scatter_tensor(o, i_, R.SUM)
            w.flush(); torch.cuda.synchronize()
            tot = sum(f.double() for f in full)
            for r, o in enumerate(outs):
                assert torch.allclose(o.double(), tot[r * n:(r + 1) * n], rtol=1e-6, atol=1e-5), ("reduce_scatter", n, r)
            dst = world - 1
            red = [f.clone() for f in full]
            for be, t in zip(w.ranks, red):
                be.reduce(t, dst, R.SUM)
            w.flush(); torch.cuda.synchronize()
            assert torch.allclose(red[dst].double(), tot, rtol=1e-6, atol=1e-5), ("reduce", n)
            # gather to / scatter from root 1 through the C ABI (flat [W][n] buffers)
            flat = torch.full((world * n,), -1.0, device=dev)
            for r, be in enumerate(w.ranks):
                assert lib.dcp_all_gather(be.handle, ins[r].data_ptr(), flat.data_ptr() if r == 1 else None, n * 4, 1, CH_USER, CTAS, st) == 0
            w.flush(); torch.cuda.synchronize()
            assert torch.equal(flat, exp), ("gather", n)
            outs = [torch.empty(n, device=dev) for _ in range(world)]
            for r, be in enumerate(w.ranks):
                assert lib.dcp_scatter(be.handle, exp.data_ptr() if r == 1 else None, outs[r].data_ptr(), n * 4, 1, CH_USER, CTAS, st) == 0
            w.flush(); torch.cuda.synchronize()
            assert all(torch.equal(o, ins[r]) for r, o in enumerate(outs)), ("scatter", n)
        w.check_health()


def test_send_recv_barrier_and_mixed_sequences():
    """Point-to-point pipelining (messages larger than a half slot), and a stress of DIFFERENT staged collectives back to back on
    one channel: all of them pick their bounce half from the same launch counter, so no kind may overwrite a half another kind's
    peers are still reading."""
    world = 4
    dev = torch.device("cuda", 0)
    with EmulatedWorld(world, staging_mb=8, p2p_slot_mb=2) as w:
        for n in (0, 1, 5000, (3 << 20) // 4 + 11):          # the last one needs 3+ half-slot chunks
            src = [_pattern(max(n, 1), r, torch.float32, dev)[:n] for r in range(world)]
            dst = [torch.full((n,), -7.0, device=dev) for _ in range(world)]
            # ring: r sends to r+1, then receives from r-1 (the order a real program would use on rank r)
            for r, be in enumerate(w.ranks):
                if r % 2 == 0:
                    be.send(src[r], (r + 1) % world); be.recv(dst[r], (r - 1) % world)
                else:
                    be.recv(dst[r], (r - 1) % world); be.send(src[r], (r + 1) % world)
            w.flush(); torch.cuda.synchronize(); w.check_health()
            for r in range(world):
                assert torch.equal(dst[r], src[(r - 1) % world]), ("send/recv", n, r)
        for be in w.ranks:
            be.barrier()
        w.flush(); torch.cuda.synchronize(); w.check_health()
        # mixed staged sequence, repeated
        n = 70001
        for it in range(4):
            a = [_pattern(n, r + it, torch.float32, dev) for r in range(world)]
            g_in = [_pattern(n, 10 * r + it, torch.float32, dev) for r in range(world)]
            g_out = [torch.empty(world * n, device=dev) for _ in range(world)]
            rs_in = [_pattern(world * n, r - it, torch.float32, dev) for r in range(world)]
            rs_out = [torch.empty(n, device=dev) for _ in range(world)]
            small = [torch.full((64,), float(r + it), device=dev) for r in range(world)]
            b = [_pattern(n, 100 + r, torch.float32, dev) for r in range(world)]
            for r, be in enumerate(w.ranks):
                be.reduce_scatter_tensor(rs_out[r], rs_in[r], R.SUM)
                be.all_gather_into_tensor(g_out[r], g_in[r])
                be.all_reduce(small[r], R.SUM, algo=1, max_ctas=CTAS)
                be.all_reduce(a[r], R.SUM, algo=2, max_ctas=CTAS)
                be.broadcast(b[r], it % world)
            w.flush(); torch.cuda.synchronize(); w.check_health()
            tot = sum(x.double() for x in rs_in)
            exp_a = sum(_pattern(n, r + it, torch.float64, dev) for r in range(world))
            for r in range(world):
                assert torch.allclose(rs_out[r].double(), tot[r * n:(r + 1) * n], rtol=1e-6, atol=1e-5), ("mixed rs", it, r)
                assert torch.equal(g_out[r], torch.cat(g_in)), ("mixed ag", it, r)
                assert torch.allclose(small[r], torch.full_like(small[r], float(sum(q + it for q in range(world))))), ("mixed one-shot", it, r)
                assert torch.allclose(a[r].double(), exp_a, rtol=1e-6, atol=1e-6), ("mixed two-shot", it, r)
                assert torch.equal(b[r], _pattern(n, 100 + it % world, torch.float32, dev)), ("mixed broadcast", it, r)


def test_mismatched_collectives_are_refused():
    with EmulatedWorld(2) as w:
        dev = torch.device("cuda", 0)
        a, b = torch.ones(100, device=dev), torch.ones(100, device=dev)
        w.ranks[0].all_reduce(a, R.SUM, algo=2, max_ctas=CTAS)
        w.ranks[1].broadcast(b, 0)
        with pytest.raises(RuntimeError, match="different collectives"):
            w.flush()
        w.ranks[0].all_reduce(a, R.SUM, algo=2, max_ctas=CTAS)        # only one rank shows up
        with pytest.raises(RuntimeError, match="issued"):
            w.flush()


@pytest.mark.parametrize("kind", [0, 1, 2])
def test_fused_reduce_scatter_optimizer_all_gather(kind):
    """dcp_fused_rs_opt_ag (ZeRO-1-shaped step: pull-reduce own shard, update, push parameters) on 4 emulated ranks vs the same
    update computed in fp64 on the full, averaged gradient."""
    world, n = 4, 100003
    dev = torch.device("cuda", 0)
    with EmulatedWorld(world) as w:
        lib = w.lib
        lib.dcp_fused_shard_elems.restype = C.c_size_t; lib.dcp_fused_shard_elems.argtypes = [C.c_size_t, C.c_int, C.c_int]
        lib.dcp_fused_rs_opt_ag.restype = C.c_int
        lib.dcp_fused_rs_opt_ag.argtypes = [C.c_void_p] * 5 + [C.c_size_t, C.c_void_p] + [C.c_int] * 5 + [C.c_void_p]
        shard = lib.dcp_fused_shard_elems(n, world, 0)
        npad = (n + 7) // 8 * 8
        gen = torch.Generator(device=dev).manual_seed(5)
        p0 = torch.randn(n, device=dev, generator=gen)
        grads = [torch.randn(n, device=dev, generator=gen) for _ in range(world)]
        gs, ps, s1, s2 = [], [], [], []
        for be, g in zip(w.ranks, grads):
            gt = be.empty_symmetric(npad, torch.float32); gt.zero_(); gt[:n].copy_(g); gs.append(gt)
            pt = be.empty_symmetric(npad, torch.float32); pt.zero_(); pt[:n].copy_(p0); ps.append(pt)
            s1.append(torch.zeros(shard, device=dev)); s2.append(torch.zeros(shard, device=dev))
        # hyper-parameter block: [lr, h1..h5, step, grad_scale]
        hyper = {0: [0.5, 0.9, 1e-6, 0.0, 0.0, 0.0], 1: [0.1, 0.9, 0.0, 0.0, 0.0, 0.0], 2: [1e-2, 0.9, 0.999, 1e-8, 0.01, 1.0]}[kind]
        hp = torch.tensor(hyper + [0.0, 1.0], device=dev)
        ref_p = p0.double().clone(); r1 = torch.zeros(n, dtype=torch.float64, device=dev); r2 = torch.zeros_like(r1)
        gavg = sum(g.double() for g in grads) / world
        st = torch.cuda.current_stream().cuda_stream
        for step in (1, 2):
            hp[6] = float(step)
            for r, be in enumerate(w.ranks):
                rc = lib.dcp_fused_rs_opt_ag(be.handle, gs[r].data_ptr(), ps[r].data_ptr(), s1[r].data_ptr(), s2[r].data_ptr(), n, hp.data_ptr(),
                                             kind, 0, 0, CH_OPT, CTAS, st)
                assert rc == 0, lib.dcp_last_error()
            w.flush(); torch.cuda.synchronize(); w.check_health()
            g = gavg
            if kind == 0:
                lr, rho, eps = hyper[0], hyper[1], hyper[2]
                r1 = rho * r1 + (1 - rho) * g * g
                delta = torch.sqrt(r2 + eps) / torch.sqrt(r1 + eps) * g
                r2 = rho * r2 + (1 - rho) * delta * delta
                ref_p = ref_p - lr * delta
            elif kind == 1:
                lr, mom = hyper[0], hyper[1]
                r1 = g.clone() if step == 1 else mom * r1 + g
                ref_p = ref_p - lr * r1
            else:
                lr, b1, b2, eps, wd = hyper[0], hyper[1], hyper[2], hyper[3], hyper[4]
                ref_p = ref_p * (1 - lr * wd)
                r1 = b1 * r1 + (1 - b1) * g; r2 = b2 * r2 + (1 - b2) * g * g
                ref_p = ref_p - lr / (1 - b1 ** step) * r1 / (torch.sqrt(r2 / (1 - b2 ** step)) + eps)
            for r in range(world):
                assert torch.allclose(ps[r][:n].double(), ref_p, rtol=2e-4, atol=2e-5), (kind, step, r, (ps[r][:n].double() - ref_p).abs().max().item())
            assert all(torch.equal(ps[0], q) for q in ps[1:]), "replicas differ after the fused step"
