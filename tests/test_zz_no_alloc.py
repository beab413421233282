"""The allocation-free BatchNorm forward (dcp_bn_forward_noalloc) called straight through the C ABI.

It serves hosts that cannot hand the library scratch memory (no PyTorch allocator behind a C/C++ caller, or a
process whose own arena owns the whole GPU, so cudaMalloc would fail and stall the device).  The scratch is a
32 KiB __device__ array inside libdcp_cuda.so that every call shares, so the things worth checking are:
same numbers as the scratch-taking entry points (dcp_bn_stats + dcp_bn_apply), also on a SECOND call right after the
first (the array still holds the first call's sums and must be cleared), and a clear refusal above the size limit.
This file sorts last on purpose: if it fails under `pytest -x`, every other test has already run.
"""
import ctypes as C

import pytest
import torch

from distributed_compute_pytorch_b200 import _native

pytestmark = pytest.mark.gpu
vp, i, l, f, d = C.c_void_p, C.c_int, C.c_long, C.c_float, C.c_double


def _abi():
    lib = C.CDLL(str(_native.cuda_lib_path()))
    lib.dcp_last_error.restype = C.c_char_p
    lib.dcp_bn_forward_noalloc.restype = i
    lib.dcp_bn_forward_noalloc.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp, vp, l, i, f, f, i, i, vp]
    lib.dcp_bn_stats.restype = i
    lib.dcp_bn_stats.argtypes = [vp, vp, l, i, i, vp]
    lib.dcp_bn_apply.restype = i
    lib.dcp_bn_apply.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, l, i, d, f, f, i, i, i, vp]
    return lib


def _run_pair(lib, M, Cc, relu, seed, dtype=torch.float32, device=0):
    """Returns (outputs of the existing entry points, outputs of the no-alloc entry point) for one input."""
    dev = torch.device("cuda", device)
    code = {torch.float32: 0, torch.bfloat16: 1}[dtype]
    g = torch.Generator(device=dev).manual_seed(seed)
    x = (torch.randn(M, Cc, device=dev, generator=g) * 3 + 1.5).to(dtype)
    gamma = torch.rand(Cc, device=dev, generator=g) + 0.5
    beta = torch.randn(Cc, device=dev, generator=g)
    stream = torch.cuda.current_stream(dev).cuda_stream
    outs = []
    for which in ("existing", "noalloc"):
        y = torch.empty_like(x)
        rm, rv = torch.zeros(Cc, device=dev), torch.ones(Cc, device=dev)
        nbt = torch.zeros((), dtype=torch.int64, device=dev)
        mean, rstd = torch.empty(Cc, device=dev), torch.empty(Cc, device=dev)
        if which == "existing":
            scratch = torch.empty(4 * Cc, dtype=torch.float64, device=dev)   # the allocation the new entry point avoids
            assert lib.dcp_bn_stats(x.data_ptr(), scratch.data_ptr(), M, Cc, code, stream) == 0
            assert lib.dcp_bn_apply(x.data_ptr(), y.data_ptr(), scratch.data_ptr(), gamma.data_ptr(), beta.data_ptr(), rm.data_ptr(),
                                    rv.data_ptr(), nbt.data_ptr(), mean.data_ptr(), rstd.data_ptr(), M, Cc, float(M), 1e-5, 0.1, relu, 1, code,
                                    stream) == 0
        else:
            rc = lib.dcp_bn_forward_noalloc(x.data_ptr(), y.data_ptr(), gamma.data_ptr(), beta.data_ptr(), rm.data_ptr(), rv.data_ptr(),
                                            nbt.data_ptr(), mean.data_ptr(), rstd.data_ptr(), M, Cc, 1e-5, 0.1, relu, code, stream)
            assert rc == 0, lib.dcp_last_error()
        torch.cuda.synchronize(dev)
        outs.append((y, mean, rstd, rm, rv, nbt))
    return outs


def test_noalloc_matches_existing_entry_point_on_two_calls_in_a_row():
    lib = _abi()
    # first call (also performs the per-device address look-up), then immediately a second one with other data and
    # a different channel count, so stale sums from call 1 would show up if the array were not cleared
    for M, Cc, relu, seed in ((4096, 512, 1, 1), (128, 128, 0, 2)):
        ref, got = _run_pair(lib, M, Cc, relu, seed)
        for a, b, name in zip(got, ref, ("y", "mean", "rstd", "running_mean", "running_var", "num_batches_tracked")):
            if a.dtype == torch.int64:
                assert a.item() == b.item() == 1, name
            else:
                assert torch.allclose(a, b, rtol=1e-4, atol=1e-4), f"{name} differs (M={M}, C={Cc}): {(a - b).abs().max().item():.3e}"
        # and against the plain PyTorch fp32 formula
        g = torch.Generator(device="cuda").manual_seed(seed)
        x = torch.randn(M, Cc, device="cuda", generator=g) * 3 + 1.5
        assert torch.allclose(got[1], x.mean(0), rtol=1e-4, atol=1e-4)
        assert torch.allclose(got[2], torch.rsqrt(x.var(0, unbiased=False) + 1e-5), rtol=1e-4, atol=1e-4)


def test_noalloc_bf16_activations_two_calls_in_a_row():
    """bf16 in / out (statistics and running buffers stay fp32): the entry point dispatches on the dtype code like the existing
    ones, so the two must agree bit for bit on y and to rounding on the statistics - again on a second call right after the first."""
    lib = _abi()
    for M, Cc, relu, seed in ((2048, 256, 1, 3), (96, 64, 0, 4)):
        ref, got = _run_pair(lib, M, Cc, relu, seed, dtype=torch.bfloat16)
        for a, b, name in zip(got, ref, ("y", "mean", "rstd", "running_mean", "running_var", "num_batches_tracked")):
            if a.dtype == torch.int64:
                assert a.item() == b.item() == 1, name
            elif a.dtype == torch.bfloat16:
                assert torch.allclose(a.float(), b.float(), rtol=2e-2, atol=2e-2), f"{name} differs (bf16, M={M}, C={Cc})"
            else:
                assert torch.allclose(a, b, rtol=1e-4, atol=1e-4), f"{name} differs (bf16, M={M}, C={Cc}): {(a - b).abs().max().item():.3e}"


@pytest.mark.multigpu
def test_noalloc_second_device_has_its_own_array():
    """Every GPU that loads the library holds its own copy of the scratch array; its address is looked up per device, with that
    device current.  A call on device 1 right after calls on device 0 must not use device 0's address."""
    lib = _abi()
    _run_pair(lib, 512, 128, 1, 5, device=0)
    with torch.cuda.device(1):
        ref, got = _run_pair(lib, 1024, 320, 1, 6, device=1)
    for a, b, name in zip(got, ref, ("y", "mean", "rstd", "running_mean", "running_var", "num_batches_tracked")):
        if a.dtype == torch.int64:
            assert a.item() == b.item() == 1, name
        else:
            assert torch.allclose(a, b, rtol=1e-4, atol=1e-4), f"{name} differs on device 1: {(a - b).abs().max().item():.3e}"


def test_noalloc_rejects_sizes_beyond_the_built_in_scratch():
    lib = _abi()
    Cc = 2048 + 32
    x = torch.zeros(8, Cc, device="cuda"); y = torch.empty_like(x)
    m = torch.empty(Cc, device="cuda"); r = torch.empty(Cc, device="cuda")
    rc = lib.dcp_bn_forward_noalloc(x.data_ptr(), y.data_ptr(), None, None, None, None, None, m.data_ptr(), r.data_ptr(), 8, Cc, 1e-5, 0.1, 0, 0,
                                    torch.cuda.current_stream().cuda_stream)
    assert rc == -1
    msg = lib.dcp_last_error().decode()
    assert "2048" in msg and "exceeds" in msg
    # the limit itself is accepted
    Cc = 2048
    x = torch.randn(64, Cc, device="cuda"); y = torch.empty_like(x)
    m = torch.empty(Cc, device="cuda"); r = torch.empty(Cc, device="cuda")
    assert lib.dcp_bn_forward_noalloc(x.data_ptr(), y.data_ptr(), None, None, None, None, None, m.data_ptr(), r.data_ptr(), 64, Cc, 1e-5, 0.1, 0, 0,
                                      torch.cuda.current_stream().cuda_stream) == 0
    torch.cuda.synchronize()
    assert torch.allclose(m, x.mean(0), rtol=1e-4, atol=1e-4)
