"""tcgen05 / TMEM / TMA tensor-core paths (csrc/ops/gemm_tc.cu): shape gating + launchers.

``*_supported`` decide per shape whether the tensor-core kernel can take the op; everything else
falls back to the CUDA-core kernels in conv_linear_simt.cu.  fp32 tensors are multiplied in tf32 with
fp32 accumulation (what cuDNN's default does for the reference's convolutions); bf16 stays bf16.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

_lib = None
NUM_SMS = 148


def declare(lib) -> None:
    global _lib
    vp, i, l = C.c_void_p, C.c_int, C.c_long
    lib.dcp_tc_gemm.restype = C.c_int
    lib.dcp_tc_gemm.argtypes = [vp, vp, vp, vp, i, i, i, l, l, l, i, i, i, i, i, i, i, vp]
    lib.dcp_tc_gemm_ex.restype = C.c_int
    lib.dcp_tc_gemm_ex.argtypes = [vp, vp, vp, vp, i, i, i, l, l, l, i, i, i, i, i, i, i, i, vp]
    lib.dcp_tc_conv.restype = C.c_int
    lib.dcp_tc_conv.argtypes = [vp, vp, vp, vp] + [i] * 17 + [vp]
    lib.dcp_tc_conv_strided.restype = C.c_int
    lib.dcp_tc_conv_strided.argtypes = [vp, vp, vp, vp] + [i] * 18 + [vp]
    lib.dcp_tc_conv_wgrad_strided.restype = C.c_int
    lib.dcp_tc_conv_wgrad_strided.argtypes = [vp, vp, vp] + [i] * 16 + [vp]
    lib.dcp_gemm_set_workspace.restype = C.c_int
    lib.dcp_gemm_set_workspace.argtypes = [vp, C.c_size_t]
    lib.dcp_conv_bf16_persistent.restype = C.c_int
    lib.dcp_conv_bf16_persistent.argtypes = [vp, vp, vp, vp] + [i] * 17 + [vp]
    lib.dcp_upsample2_zero.restype = C.c_int
    lib.dcp_upsample2_zero.argtypes = [vp, vp] + [i] * 7 + [vp]
    lib.dcp_im2col_small_c.restype = C.c_int
    lib.dcp_im2col_small_c.argtypes = [vp, vp] + [i] * 12 + [vp]
    lib.dcp_tc_conv_wgrad.restype = C.c_int
    lib.dcp_tc_conv_wgrad.argtypes = [vp, vp, vp] + [i] * 15 + [vp]
    lib.dcp_splitk_reduce.restype = C.c_int
    lib.dcp_splitk_reduce.argtypes = [vp, vp, vp, i, l, i, vp]
    lib.dcp_gemm_bf16_persistent.restype = C.c_int
    lib.dcp_gemm_bf16_persistent.argtypes = [vp, vp, vp, vp, i, i, i, l, l, l, i, i, i, i, i, i, i, vp]
    lib.dcp_splitk_reduce2.restype = C.c_int
    lib.dcp_splitk_reduce2.argtypes = [vp, vp, vp, l, i, l, vp]
    lib.dcp_flip_weights.restype = C.c_int
    lib.dcp_flip_weights.argtypes = [vp, vp, i, i, i, i, i, vp]
    _lib = lib


def _k():
    from . import cuda
    if _lib is None:
        cuda.lib()          # loads libdcp_cuda.so and calls declare() above
    return cuda


def _bk(dtype) -> int:
    return 32 if dtype == torch.float32 else 64


def _ok_dtype(dtype) -> bool:
    return dtype in (torch.float32, torch.bfloat16)


# fp32 Linear layers default to 3xTF32 (fp32-class accuracy, like the reference's cuBLAS fp32 matmul);
# DCP_LINEAR_PRECISION=tf32 selects the single-pass tf32 multiply instead.
PERSISTENT_CONV = os.environ.get("DCP_CONV_PERSISTENT", "1") == "1"     # bf16 implicit-GEMM convs on gemm_bf16_persistent.cu
LINEAR_X3 = os.environ.get("DCP_LINEAR_PRECISION", "tf32x3") == "tf32x3"


_workspace = None


def _ensure_workspace(device) -> None:
    """20 MB of scratch for the persistent GEMM's tail split (fp32 partial tiles of the last, partial wave); registered once."""
    global _workspace
    if _workspace is None and os.environ.get("DCP_GEMM_TAIL_SPLIT", "1") == "1":
        _workspace = torch.empty(20 << 20, dtype=torch.uint8, device=device)
        _lib.dcp_gemm_set_workspace(_workspace.data_ptr(), _workspace.numel())


def gemm(A, B, out, bias, M, N, K, lda, ldb, ldc, a_mn, b_mn, relu=False, ksplit=1, x3=None) -> None:
    """out[M,N] (+)= op(A) op(B)^T on tcgen05 (see dcp_tc_gemm_ex)."""
    k = _k()
    if A.dtype == torch.bfloat16 and _workspace is None:
        _ensure_workspace(A.device)
    x3 = LINEAR_X3 if x3 is None else x3
    k._chk(_lib.dcp_tc_gemm_ex(A.data_ptr(), B.data_ptr(), out.data_ptr(), k._p(bias), M, N, K, lda, ldb, ldc, int(a_mn), int(b_mn), k._dt(A),
                               int(out.dtype == torch.bfloat16), int(bias is not None and bias.dtype == torch.bfloat16), int(relu), int(ksplit),
                               int(bool(x3) and A.dtype == torch.float32), k._stream()))


def _ksplit(M: int, N: int, K: int, dtype) -> int:
    """K splits for skinny problems: enough CTAs to pull the operands at full rate, >= 4 K slabs per CTA."""
    if dtype == torch.bfloat16:
        # persistent kernel: 128 x 256 tiles run the tensor pipe at full rate, so fill the SMs with K slices of wide
        # tiles rather than with narrow tiles (>= 8 slabs of 64 per slice)
        tiles = -(-M // 128) * -(-N // (256 if N > 128 else 128))
        stages = K // 64
        if tiles >= 100 or stages < 16:
            return 1
        return max(1, min(stages // 8, NUM_SMS // tiles))
    tiles = -(-M // 128) * -(-N // (32 if N <= 32 else 64 if N <= 64 else 128))
    stages = K // _bk(dtype)
    if tiles >= NUM_SMS // 2 or stages < 8:
        return 1
    return max(1, min(stages // 4, NUM_SMS // tiles))


def gemm_splitk(A, B, out, bias, M, N, K, lda, ldb, ldc, a_mn, b_mn, ks, x3=None) -> None:
    """Split-K GEMM without atomics: partial tiles go to a scratch buffer with plain stores, then one reduce kernel
    sums them (and adds the bias) into ``out`` (fp32, ldc == N)."""
    k = _k()
    scratch = torch.empty(ks * M * N, dtype=torch.float32, device=out.device)
    gemm(A, B, scratch, None, M, N, K, lda, ldb, N, a_mn, b_mn, False, -ks, x3)
    k._chk(_lib.dcp_splitk_reduce(scratch.data_ptr(), out.data_ptr(), k._p(bias), ks, M * N, N, k._stream()))


def linear_supported(M: int, N: int, K: int, dtype) -> bool:
    """GEMM with output [M, N] and contraction K."""
    return _ok_dtype(dtype) and K % _bk(dtype) == 0 and M % 8 == 0 and N % 8 == 0 and M >= 8 and N >= 16


def linear_fwd(x2, w, b, y, relu) -> None:
    M, K = x2.shape
    N = w.shape[0]
    ks = _ksplit(M, N, K, x2.dtype) if (y.dtype == torch.float32 and not relu) else 1
    if ks > 1:
        gemm_splitk(x2, w, y, b, M, N, K, K, K, N, 0, 0, ks)
    else:
        gemm(x2, w, y, b, M, N, K, K, K, N, 0, 0, relu, 1)


def linear_dgrad(dy, w, dx) -> None:
    """dx[M,K] = dy[M,N] . w[N,K]: A = dy (K-major over N), B = w read MN-major."""
    M, N = dy.shape
    K = w.shape[1]
    ks = _ksplit(M, K, N, dy.dtype) if dx.dtype == torch.float32 else 1
    if ks > 1:
        gemm_splitk(dy, w, dx, None, M, K, N, N, K, K, 0, 1, ks)
    else:
        gemm(dy, w, dx, None, M, K, N, N, K, K, 0, 1, False, 1)


def linear_wgrad(dy, x2, dw32, db32) -> None:
    """dw[N,K] = dy[M,N]^T . x[M,K]: both operands MN-major (contraction over the batch M)."""
    M, N = dy.shape
    K = x2.shape[1]
    ks = _ksplit(N, K, M, dy.dtype)
    if ks > 1:
        gemm_splitk(dy, x2, dw32, None, N, K, M, N, K, K, 1, 1, ks)
    else:
        gemm(dy, x2, dw32, None, N, K, M, N, K, K, 1, 1, False, 1)
    if db32 is not None:
        k = _k()
        k._chk(k.lib().dcp_colsum(dy.data_ptr(), db32.data_ptr(), M, N, k._dt(dy), k._stream()))


def _pick_box(Wo: int, Ho: int, N: int, pixels: int = 128):
    best = None
    for wb in (1, 2, 4, 8, 16, 32, 64, 128):
        for hb in (1, 2, 4, 8, 16, 32, 64, 128):
            if pixels % (wb * hb):
                continue
            nb = pixels // (wb * hb)
            if nb < 1:
                continue
            cover = (-(-Wo // wb) * wb) * (-(-Ho // hb) * hb) * (-(-N // nb) * nb)
            key = (cover, -wb)
            if best is None or key < best[0]:
                best = (key, (wb, hb, nb))
    return best[1]


def _small_c(p, dtype) -> bool:
    """Few input channels (the 7x7 / 3-channel stem): patch matrix (im2col, zero padded to the K slab) + GEMM."""
    return _ok_dtype(dtype) and p.C < 8 and p.K % 8 == 0 and p.K >= 16 and (p.N * p.Ho * p.Wo) % 8 == 0 and os.environ.get("DCP_TC_CONV", "1") == "1"


def _kpad(p, dtype) -> int:
    bk = _bk(dtype)
    return -(-(p.R * p.S * p.C) // bk) * bk


def _patches(x, p):
    k = _k()
    kp = _kpad(p, x.dtype)
    out = torch.empty(p.N * p.Ho * p.Wo, kp, dtype=x.dtype, device=x.device)
    k._chk(_lib.dcp_im2col_small_c(x.data_ptr(), out.data_ptr(), p.N, p.H, p.W, p.C, p.R, p.S, p.stride, p.pad, p.Ho, p.Wo, kp, k._dt(x), k._stream()))
    return out, kp


def conv_fwd_supported(p, dtype) -> bool:
    if _small_c(p, dtype):
        return True
    return (_ok_dtype(dtype) and p.stride in (1, 2) and p.C % _bk(dtype) == 0 and p.K % 8 == 0 and p.K >= 16
            and os.environ.get("DCP_TC_CONV", "1") == "1")


def _is_1x1_gemm(p, dtype) -> bool:
    """1x1 / stride 1 / no padding on channels-last bf16 tensors IS a GEMM over the flattened pixels: it goes to the
    persistent bf16 GEMM (gemm_bf16_persistent.cu) instead of the implicit-conv kernel."""
    return p.R == 1 and p.S == 1 and p.pad == 0 and p.stride == 1 and dtype == torch.bfloat16


def conv2d_fwd(x, w, b, y, p, relu) -> None:
    k = _k()
    if _is_1x1_gemm(p, x.dtype) and p.C % 64 == 0:
        gemm(x, w, y, b, p.N * p.H * p.W, p.K, p.C, p.C, p.C, p.K, 0, 0, relu, 1)
        return
    if _small_c(p, x.dtype):
        pm, kp = _patches(x, p)
        kd = p.R * p.S * p.C
        wp = torch.zeros(p.K, kp, dtype=w.dtype, device=w.device)
        wp[:, :kd] = w.permute(0, 2, 3, 1).reshape(p.K, kd)              # physical (K, R, S, C) order, zero padded columns
        gemm(pm, wp, y, b, p.N * p.Ho * p.Wo, p.K, kp, kp, kp, p.K, 0, 0, relu, 1, x3=False)
        return
    wb, hb, nb = _pick_box(p.Wo, p.Ho, p.N)
    if x.dtype == torch.bfloat16 and PERSISTENT_CONV:
        k._chk(_lib.dcp_conv_bf16_persistent(x.data_ptr(), w.data_ptr(), y.data_ptr(), k._p(b), p.N, p.H, p.W, p.C, p.K, p.R, p.S, p.pad, p.stride, p.Ho, p.Wo,
                                             wb, hb, nb, int(y.dtype == torch.bfloat16), int(b is not None and b.dtype == torch.bfloat16), int(relu), k._stream()))
        return
    k._chk(_lib.dcp_tc_conv_strided(x.data_ptr(), w.data_ptr(), y.data_ptr(), k._p(b), p.N, p.H, p.W, p.C, p.K, p.R, p.S, p.pad, p.stride, p.Ho, p.Wo,
                                    wb, hb, nb, k._dt(x), int(y.dtype == torch.bfloat16), int(b is not None and b.dtype == torch.bfloat16), int(relu),
                                    k._stream()))


def conv_dgrad_supported(p, dtype) -> bool:
    return (_ok_dtype(dtype) and p.stride in (1, 2) and p.K % _bk(dtype) == 0 and p.C % 8 == 0 and p.C >= 16 and p.R - 1 - p.pad >= 0
            and os.environ.get("DCP_TC_CONV", "1") == "1")


def conv2d_dgrad(dy, w, dx, p) -> None:
    """dx = conv(dy, flipped weights) with padding R-1-pad: the same implicit-GEMM kernel as forward."""
    k = _k()
    if p.stride == 2:
        # stride 2: interleave dy with zeros (one small kernel), then it is the stride-1 data-gradient of that tensor
        hup, wup = p.H + 2 * p.pad - p.R + 1, p.W + 2 * p.pad - p.S + 1
        up = torch.empty(p.N * hup * wup * p.K, dtype=dy.dtype, device=dy.device)
        k._chk(_lib.dcp_upsample2_zero(dy.data_ptr(), up.data_ptr(), p.N, p.Ho, p.Wo, hup, wup, p.K, k._dt(dy), k._stream()))
        p1 = type(p)(p.N, p.H, p.W, p.C, p.K, p.R, p.S, 1, p.pad, hup, wup)
        conv2d_dgrad(up, w, dx, p1)
        return
    if _is_1x1_gemm(p, dy.dtype) and p.K % 64 == 0:
        gemm(dy, w, dx, None, p.N * p.H * p.W, p.C, p.K, p.K, p.C, p.C, 0, 1, False, 1)      # dx[M,C] = dy[M,K] . w[K,C]
        return
    wt = torch.empty(p.C * p.R * p.S * p.K, dtype=w.dtype, device=w.device)
    k._chk(_lib.dcp_flip_weights(w.data_ptr(), wt.data_ptr(), p.K, p.R, p.S, p.C, k._dt(w), k._stream()))
    wb, hb, nb = _pick_box(p.W, p.H, p.N)
    if dy.dtype == torch.bfloat16 and PERSISTENT_CONV:
        k._chk(_lib.dcp_conv_bf16_persistent(dy.data_ptr(), wt.data_ptr(), dx.data_ptr(), None, p.N, p.Ho, p.Wo, p.K, p.C, p.R, p.S, p.R - 1 - p.pad, 1, p.H, p.W,
                                             wb, hb, nb, int(dx.dtype == torch.bfloat16), 0, 0, k._stream()))
        return
    k._chk(_lib.dcp_tc_conv(dy.data_ptr(), wt.data_ptr(), dx.data_ptr(), None, p.N, p.Ho, p.Wo, p.K, p.C, p.R, p.S, p.R - 1 - p.pad, p.H, p.W,
                            wb, hb, nb, k._dt(dy), int(dx.dtype == torch.bfloat16), 0, 0, k._stream()))


def conv_wgrad_launch(dy, x, dw32, N, H, W, Cc, Kout, R, S, pad, Ho, Wo, wb, hb, nb, zsplit, stride=1) -> None:
    """Weight-gradient kernel; with several pixel splits the partial filters meet in a scratch buffer + reduce."""
    k = _k()
    if zsplit > 1:
        n = Kout * R * S * Cc
        scratch = torch.empty(zsplit * n, dtype=torch.float32, device=dw32.device)
        k._chk(_lib.dcp_tc_conv_wgrad_strided(dy.data_ptr(), x.data_ptr(), scratch.data_ptr(), N, H, W, Cc, Kout, R, S, pad, stride, Ho, Wo, wb, hb, nb,
                                              k._dt(dy), -zsplit, k._stream()))
        k._chk(_lib.dcp_splitk_reduce(scratch.data_ptr(), dw32.data_ptr(), None, zsplit, n, 4, k._stream()))
    else:
        k._chk(_lib.dcp_tc_conv_wgrad_strided(dy.data_ptr(), x.data_ptr(), dw32.data_ptr(), N, H, W, Cc, Kout, R, S, pad, stride, Ho, Wo, wb, hb, nb,
                                              k._dt(dy), 1, k._stream()))


def conv_wgrad_supported(p, dtype) -> bool:
    if _small_c(p, dtype) and (p.N * p.Ho * p.Wo) % _bk(dtype) == 0:
        return True
    return (_ok_dtype(dtype) and p.stride in (1, 2) and p.C % _bk(dtype) == 0 and p.K % 8 == 0 and p.K >= 16
            and os.environ.get("DCP_TC_CONV", "1") == "1")


def conv2d_wgrad(dy, x, dw32, db32, p) -> None:
    """dw (fp32 [K,R,S,C]) on tcgen05 with both operands read MN-major from the channels-last tensors."""
    k = _k()
    M = p.N * p.H * p.W
    if _is_1x1_gemm(p, dy.dtype) and M % 64 == 0:
        ks = _ksplit(p.K, p.C, M, dy.dtype)                                                  # dw[K,C] = dy[M,K]^T . x[M,C]
        if ks > 1:
            gemm_splitk(dy, x, dw32, None, p.K, p.C, M, p.K, p.C, p.C, 1, 1, ks)
        else:
            gemm(dy, x, dw32, None, p.K, p.C, M, p.K, p.C, p.C, 1, 1, False, 1)
        if db32 is not None:
            k._chk(k.lib().dcp_colsum(dy.data_ptr(), db32.data_ptr(), M, p.K, k._dt(dy), k._stream()))
        return
    if _small_c(p, dy.dtype):
        # stem: dw[K, Kpad] = dy[M, K]^T . patches[M, Kpad] (patch matrix recomputed: cheaper than keeping 300 MB alive)
        pm, kp = _patches(x, p)
        Mo, kd = p.N * p.Ho * p.Wo, p.R * p.S * p.C
        dwp = torch.empty(p.K, kp, dtype=torch.float32, device=dy.device)
        ks = _ksplit(p.K, kp, Mo, dy.dtype)
        if ks > 1:
            gemm_splitk(dy, pm, dwp, None, p.K, kp, Mo, p.K, kp, kp, 1, 1, ks, x3=False)
        else:
            gemm(dy, pm, dwp, None, p.K, kp, Mo, p.K, kp, kp, 1, 1, False, 1, x3=False)
        dw32.permute(0, 2, 3, 1).copy_(dwp[:, :kd].view(p.K, p.R, p.S, p.C))
        if db32 is not None:
            k._chk(k.lib().dcp_colsum(dy.data_ptr(), db32.data_ptr(), Mo, p.K, k._dt(dy), k._stream()))
        return
    chunk = _bk(dy.dtype)
    wb, hb, nb = _pick_box(p.Wo, p.Ho, p.N, chunk)
    boxes = -(-p.Wo // wb) * -(-p.Ho // hb) * -(-p.N // nb)
    units = p.R * p.S * (p.C // chunk)
    groups_total = -(-units // (128 // chunk))
    bn = 32 if p.K <= 32 else 64 if p.K <= 64 else 128
    g = max(1, min(3, groups_total, 512 // bn))
    ctas = -(-groups_total // g) * -(-p.K // bn)
    zsplit = max(1, min(boxes, NUM_SMS // ctas if ctas < NUM_SMS else 1))
    conv_wgrad_launch(dy, x, dw32, p.N, p.H, p.W, p.C, p.K, p.R, p.S, p.pad, p.Ho, p.Wo, wb, hb, nb, zsplit, p.stride)
    if db32 is not None:
        k._chk(k.lib().dcp_colsum(dy.data_ptr(), db32.data_ptr(), p.N * p.Ho * p.Wo, p.K, k._dt(dy), k._stream()))
