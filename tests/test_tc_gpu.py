"""tcgen05 GEMM / implicit-GEMM conv kernels vs fp32 references: every operand-major combination,
tf32 and bf16, split-K, bias/ReLU epilogue, ragged M/N, conv forward + data-gradient."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _tc():
    from distributed_compute_pytorch_b200.ops import cuda, tc
    cuda.lib()
    return tc


def _rel_err(a, b):
    return ((a.double() - b.double()).abs().max() / (b.double().abs().max() + 1e-12)).item()


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("a_mn,b_mn", [(0, 0), (0, 1), (1, 1), (1, 0)])
@pytest.mark.parametrize("M,N,K", [(128, 128, 256), (256, 64, 128), (200, 72, 384), (128, 32, 64), (384, 256, 1024)])
def test_gemm_all_majors(dtype, a_mn, b_mn, M, N, K):
    tc = _tc()
    torch.manual_seed(M + N + K)
    A = torch.randn(M, K, device=DEV)
    B = torch.randn(N, K, device=DEV)
    ref = A.to(dtype).double() @ B.to(dtype).double().t()
    Ast = (A.t().contiguous() if a_mn else A).to(dtype).contiguous()
    Bst = (B.t().contiguous() if b_mn else B).to(dtype).contiguous()
    out = torch.full((M, N), float("nan"), device=DEV)
    tc.gemm(Ast, Bst, out, None, M, N, K, Ast.shape[1], Bst.shape[1], N, a_mn, b_mn, x3=False)
    torch.cuda.synchronize()
    tol = 2e-3 if dtype == torch.float32 else 1e-5   # tf32 multiplies truncate fp32 inputs; bf16 inputs are exact
    assert _rel_err(out, ref) < tol, (_rel_err(out, ref), dtype, a_mn, b_mn)


@pytest.mark.parametrize("a_mn,b_mn", [(0, 0), (0, 1), (1, 1), (1, 0)])
@pytest.mark.parametrize("M,N,K,ks", [(128, 128, 9216, 36), (200, 72, 384, 1), (128, 9216, 128, 1)])
def test_gemm_3xtf32_is_fp32_accurate(a_mn, b_mn, M, N, K, ks):
    """fp32 Linear layers use the 3xTF32 split: error must be at fp32 level, not tf32 level."""
    tc = _tc()
    torch.manual_seed(1)
    A = torch.randn(M, K, device=DEV); B = torch.randn(N, K, device=DEV)
    ref = A.double() @ B.double().t()
    Ast = (A.t().contiguous() if a_mn else A).contiguous(); Bst = (B.t().contiguous() if b_mn else B).contiguous()
    out = torch.zeros(M, N, device=DEV)
    tc.gemm(Ast, Bst, out, None, M, N, K, Ast.shape[1], Bst.shape[1], N, a_mn, b_mn, False, ks, x3=True)
    e3 = _rel_err(out, ref)
    out1 = torch.zeros(M, N, device=DEV)
    tc.gemm(Ast, Bst, out1, None, M, N, K, Ast.shape[1], Bst.shape[1], N, a_mn, b_mn, False, ks, x3=False)
    e1 = _rel_err(out1, ref)
    fp32 = _rel_err(A @ B.t(), ref)      # what cuBLAS fp32 gives
    assert e3 < 1e-5 and e3 < e1 / 50, (e3, e1, fp32)


def test_gemm_splitk_bias_relu_bf16_out():
    tc = _tc()
    torch.manual_seed(0)
    M, N, K = 128, 128, 9216   # fc1 of the reference (ref:main.py:27)
    A, B, bias = torch.randn(M, K, device=DEV), torch.randn(N, K, device=DEV) * 0.02, torch.randn(N, device=DEV)
    ref = A.double() @ B.double().t() + bias.double()
    for ks in (1, 4, 144):
        out = torch.zeros(M, N, device=DEV)
        tc.gemm(A, B, out, bias, M, N, K, K, K, N, 0, 0, False, ks, x3=False)
        assert _rel_err(out, ref) < 2e-3, ks
        if ks > 1:   # the same split through scratch + reduce (no atomics, nothing to pre-zero)
            out2 = torch.full((M, N), float("nan"), device=DEV)
            tc.gemm_splitk(A, B, out2, bias, M, N, K, K, K, N, 0, 0, ks, x3=False)
            assert _rel_err(out2, ref) < 2e-3, ks
    out = torch.empty(M, N, device=DEV)
    tc.gemm(A, B, out, bias, M, N, K, K, K, N, 0, 0, True, 1)
    assert _rel_err(out, ref.clamp(min=0)) < 2e-3
    Ab, Bb, bb = A.bfloat16(), B.bfloat16(), bias.bfloat16()
    outb = torch.empty(M, N, device=DEV, dtype=torch.bfloat16)
    tc.gemm(Ab, Bb, outb, bb, M, N, K, K, K, N, 0, 0, False, 1)
    refb = Ab.double() @ Bb.double().t() + bb.double()
    assert _rel_err(outb, refb) < 1e-2


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("N,C,H,W,K,R,pad,stride", [(128, 32, 26, 26, 64, 3, 0, 1), (4, 64, 14, 14, 128, 3, 1, 1), (6, 128, 7, 9, 64, 1, 0, 1),
                                                      (2, 64, 20, 20, 32, 3, 1, 1), (4, 64, 28, 28, 128, 3, 1, 2), (4, 128, 28, 28, 256, 1, 0, 2),
                                                      (3, 64, 15, 17, 64, 3, 1, 2), (8, 3, 32, 32, 64, 7, 3, 2)])
def test_conv_fwd_and_dgrad(dtype, N, C, H, W, K, R, pad, stride):
    """Implicit-GEMM convolution forward / data-gradient / weight-gradient on tcgen05, stride 1 and 2 (TMA elementStrides;
    zero-interleaved dy for the stride-2 data-gradient), 1x1 as plain GEMM, and the few-channel stem via patch matrix + GEMM."""
    from distributed_compute_pytorch_b200.ops import cuda
    tc = _tc()
    if C >= 8 and C % (32 if dtype == torch.float32 else 64):
        pytest.skip("C not a multiple of the K slab for this dtype")
    torch.manual_seed(0)
    x = torch.randn(N, C, H, W, device=DEV).to(dtype).contiguous(memory_format=torch.channels_last)
    w = (torch.randn(K, C, R, R, device=DEV) * 0.1).to(dtype).contiguous(memory_format=torch.channels_last)
    b = torch.randn(K, device=DEV).to(dtype)
    p = cuda._convp(x.shape, w.shape, stride, pad)
    assert tc.conv_fwd_supported(p, dtype)
    y = torch.empty((N, K, p.Ho, p.Wo), dtype=dtype, device=DEV, memory_format=torch.channels_last)
    tc.conv2d_fwd(x, w, b, y, p, True)
    ref = F.relu(F.conv2d(x.double(), w.double(), b.double(), stride, pad))
    tol = 3e-3 if dtype == torch.float32 else 1.5e-2
    assert _rel_err(y, ref) < tol, _rel_err(y, ref)
    if tc.conv_dgrad_supported(p, dtype):
        dy = torch.randn_like(y)
        dx = torch.empty_like(x)
        tc.conv2d_dgrad(dy, w, dx, p)
        xr = x.double().requires_grad_(True)
        F.conv2d(xr, w.double(), None, stride, pad).backward(dy.double())
        assert _rel_err(dx, xr.grad) < tol, _rel_err(dx, xr.grad)
    if tc.conv_wgrad_supported(p, dtype):
        dy = torch.randn_like(y)
        dw32 = torch.empty_like(w, dtype=torch.float32)
        db32 = torch.empty(K, dtype=torch.float32, device=DEV)
        tc.conv2d_wgrad(dy, x, dw32, db32, p)
        wr = w.double().requires_grad_(True)
        br = b.double().requires_grad_(True)
        F.conv2d(x.double(), wr, br, stride, pad).backward(dy.double())
        assert _rel_err(dw32, wr.grad) < tol, _rel_err(dw32, wr.grad)
        assert _rel_err(db32, br.grad) < tol


@pytest.mark.parametrize("N,Cout,W", [(4, 64, 26), (37, 64, 26), (16, 32, 26), (8, 64, 14)])
def test_conv_wgrad_flat(N, Cout, W):
    """Flat-shift weight gradient (dcp_conv_wgrad_flat_tc + dcp_splitk_reduce) vs F.conv2d's fp32 weight gradient."""
    from distributed_compute_pytorch_b200.ops import cuda
    tc = _tc()
    L = cuda.lib()
    torch.manual_seed(N * Cout)
    H = W
    x = torch.randn(N, 32, H, W, device=DEV)
    dy_valid = torch.randn(N, Cout, H - 2, W - 2, device=DEV)
    # reference in fp64 on tf32-rounded... plain fp32 inputs; tf32 multiply tolerance below
    w = torch.zeros(Cout, 32, 3, 3, device=DEV, dtype=torch.float64, requires_grad=True)
    F.conv2d(x.double(), w).backward(dy_valid.double())
    ref = w.grad.permute(0, 2, 3, 1).contiguous()                    # [K, R, S, C]
    xf = x.permute(0, 2, 3, 1).contiguous().view(N * H * W, 32)
    dyf = torch.zeros(N, H, W, Cout, device=DEV)
    dyf[:, :H - 2, :W - 2, :] = dy_valid.permute(0, 2, 3, 1)
    dyf = dyf.view(N * H * W, Cout)
    P = N * H * W
    for parts in (min(148, -(-P // 128)), 3):
        scratch = torch.full((parts * Cout * 289,), float("nan"), device=DEV)
        dw = torch.empty(Cout, 3, 3, 32, device=DEV)
        db = torch.empty(Cout, device=DEV)
        cuda._chk(L.dcp_conv_wgrad_flat_tc(xf.data_ptr(), dyf.data_ptr(), scratch.data_ptr(), P, Cout, W, parts, cuda._stream()))
        cuda._chk(tc._lib.dcp_splitk_reduce2(scratch.data_ptr(), dw.data_ptr(), db.data_ptr(), Cout * 288, parts, Cout * 289, cuda._stream()))
        torch.cuda.synchronize()
        err = ((dw.double() - ref).norm() / ref.norm()).item()
        assert err < 2e-3, (parts, err)
        db_ref = dy_valid.double().sum((0, 2, 3))            # bias gradient = column sums of dy (fp32 adds, no tensor core)
        assert ((db.double() - db_ref).norm() / db_ref.norm()).item() < 1e-5


@pytest.mark.parametrize("a_mn,b_mn", [(0, 0), (0, 1), (1, 1)])
@pytest.mark.parametrize("M,N,K,ks", [(8192, 1000, 256, 1), (2048, 2304, 768, 1), (768, 768, 4096, 4), (136, 520, 192, 1),
                                      (8192, 768, 2048, 1), (8104, 760, 1024, 1)])       # the last two: 192 tiles = 148 + a 44-tile tail that is K-split
def test_bf16_persistent_gemm(a_mn, b_mn, M, N, K, ks):
    """gemm_bf16_persistent.cu: several tiles per CTA (double-buffered TMEM accumulator), ragged M / N tiles, both tile widths,
    split-K through scratch, every operand-major combination used by Linear forward / dgrad / wgrad."""
    tc = _tc()
    torch.manual_seed(M + N + K)
    A = torch.randn(M, K, device=DEV).to(torch.bfloat16)
    B = (torch.randn(N, K, device=DEV) * 0.1).to(torch.bfloat16)
    bias = torch.randn(N, device=DEV)
    ref = A.double() @ B.double().t()
    Ast = A.t().contiguous() if a_mn else A
    Bst = B.t().contiguous() if b_mn else B
    lda = M if a_mn else K
    ldb = N if b_mn else K
    if ks > 1:
        out = torch.empty(M, N, device=DEV)
        tc.gemm_splitk(Ast, Bst, out, bias, M, N, K, lda, ldb, N, a_mn, b_mn, ks)
        ref = ref + bias.double()
        assert _rel_err(out, ref) < 2e-3
    else:
        out = torch.full((M, N), float("nan"), device=DEV, dtype=torch.bfloat16)
        tc.gemm(Ast, Bst, out, bias, M, N, K, lda, ldb, N, a_mn, b_mn, True, 1)
        ref = torch.relu(ref + bias.double())
        assert _rel_err(out, ref) < 1e-2          # bf16 output rounding
        out32 = torch.full((M, N), float("nan"), device=DEV)
        tc.gemm(Ast, Bst, out32, None, M, N, K, lda, ldb, N, a_mn, b_mn, False, 1)
        assert _rel_err(out32, A.double() @ B.double().t()) < 2e-3
