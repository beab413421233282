// Element-wise / pooling kernels: ReLU, residual add+ReLU, dropout (element and channel), max-pool,
// global average pool, eval correct-count.  These are the K2/K4-K6/K10-K11/K19 call sites of the
// reference model (ref:main.py:33-37,41-42,83-85); where a producer kernel can absorb them (conv and
// BatchNorm epilogues) the layers use the fused form instead and these only serve the stand-alone case.
// All activations are channels-last ([N,H,W,C] physical) or 2-D; f32 and bf16.
#include "ops/ops_common.cuh"

namespace {
constexpr int kT = 256;

template <typename T> __global__ void k_relu_fwd(const T* __restrict__ x, T* __restrict__ y, long n) {
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) y[i] = from_f<T>(fmaxf(to_f(x[i]), 0.f));
}
// dx = dy * (y > 0)
template <typename T> __global__ void k_relu_bwd(const T* __restrict__ y, const T* __restrict__ dy, T* __restrict__ dx, long n) {
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x)
    dx[i] = to_f(y[i]) > 0.f ? dy[i] : from_f<T>(0.f);
}
template <typename T> __global__ void k_add_relu_fwd(const T* __restrict__ a, const T* __restrict__ b, T* __restrict__ y, long n) {
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x)
    y[i] = from_f<T>(fmaxf(to_f(a[i]) + to_f(b[i]), 0.f));
}

// mode 0: element-wise; 1: per (n,c) with NHWC physical layout (inner = C, hw = H*W); 2: per (n,c) NCHW physical
__device__ __forceinline__ uint64_t drop_index(long i, int mode, long inner, long hw) {
  if (mode == 0) return (uint64_t)i;
  if (mode == 1) { long c = i % inner; long n = i / (inner * hw); return (uint64_t)(n * inner + c); }
  return (uint64_t)(i / hw);
}
// y = x * keep / (1-p).  The same call with x = dy regenerates the mask for backward.
template <typename T> __global__ void k_dropout(const T* __restrict__ x, T* __restrict__ y, long n, float p, uint64_t seed, uint64_t stream,
                                                const long long* __restrict__ step, int mode, long inner, long hw) {
  const float scale = 1.f / (1.f - p);
  const uint64_t st = stream + (step ? (uint64_t)(*step) * 0x10000ull : 0ull);
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
    float u = philox_uniform(seed, st, drop_index(i, mode, inner, hw));
    y[i] = u >= p ? from_f<T>(to_f(x[i]) * scale) : from_f<T>(0.f);
  }
}

// Max-pool over channels-last input.  arg stores the window position (kh*k+kw) of the maximum.
// out_nchw: write y (and read dy in backward) in NCHW-contiguous order so a following flatten is free.
template <typename T> __global__ void k_maxpool_fwd(const T* __restrict__ x, T* __restrict__ y, uint8_t* __restrict__ arg, int N, int H, int W, int C,
                                                    int Ho, int Wo, int k, int s, int pad, int out_nchw) {
  long total = (long)N * Ho * Wo * C;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    int c = i % C; long r = i / C; int wo = r % Wo; r /= Wo; int ho = r % Ho; int n = r / Ho;
    float best = -INFINITY; int bi = 0;
    for (int kh = 0; kh < k; ++kh) for (int kw = 0; kw < k; ++kw) {
      int h = ho * s - pad + kh, w = wo * s - pad + kw;
      if (h < 0 || h >= H || w < 0 || w >= W) continue;
      float v = to_f(x[(((long)n * H + h) * W + w) * C + c]);
      if (v > best) { best = v; bi = kh * k + kw; }
    }
    long o = out_nchw ? (((long)n * C + c) * Ho + ho) * Wo + wo : i;
    y[o] = from_f<T>(best);
    arg[i] = (uint8_t)bi;
  }
}
// Gather form: every input pixel sums the windows whose maximum it was (no atomics, no zero fill).
template <typename T> __global__ void k_maxpool_bwd(const T* __restrict__ dy, const uint8_t* __restrict__ arg, T* __restrict__ dx, int N, int H, int W,
                                                    int C, int Ho, int Wo, int k, int s, int pad, int out_nchw) {
  long total = (long)N * H * W * C;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    int c = i % C; long r = i / C; int w = r % W; r /= W; int h = r % H; int n = r / H;
    float acc = 0.f;
    int ho_lo = max(0, (h + pad - k + s) / s), ho_hi = min(Ho - 1, (h + pad) / s);
    int wo_lo = max(0, (w + pad - k + s) / s), wo_hi = min(Wo - 1, (w + pad) / s);
    for (int ho = ho_lo; ho <= ho_hi; ++ho) for (int wo = wo_lo; wo <= wo_hi; ++wo) {
      int kh = h + pad - ho * s, kw = w + pad - wo * s;
      if (kh < 0 || kh >= k || kw < 0 || kw >= k) continue;
      long oi = (((long)n * Ho + ho) * Wo + wo) * C + c;
      if (arg[oi] == kh * k + kw) acc += to_f(dy[out_nchw ? (((long)n * C + c) * Ho + ho) * Wo + wo : oi]);
    }
    dx[i] = from_f<T>(acc);
  }
}

// 16-byte variants for channels-last tensors whose C is a multiple of the vector (8 bf16 / 4 f32): one thread = one pixel x
// one channel vector, 32-bit index arithmetic, the argmax bytes of a vector written as one word.
template <typename T> struct PoolVec { static constexpr int N = 16 / sizeof(T); };
template <typename T> __device__ __forceinline__ void pool_load(const T* p, float* o) {
  alignas(16) T t[PoolVec<T>::N];
  *reinterpret_cast<uint4*>(t) = *reinterpret_cast<const uint4*>(p);
#pragma unroll
  for (int i = 0; i < PoolVec<T>::N; ++i) o[i] = to_f(t[i]);
}
template <typename T> __device__ __forceinline__ void pool_store(T* p, const float* o) {
  alignas(16) T t[PoolVec<T>::N];
#pragma unroll
  for (int i = 0; i < PoolVec<T>::N; ++i) t[i] = from_f<T>(o[i]);
  *reinterpret_cast<uint4*>(p) = *reinterpret_cast<const uint4*>(t);
}
template <typename T> __global__ void k_maxpool_fwd_vec(const T* __restrict__ x, T* __restrict__ y, uint8_t* __restrict__ arg, int N, int H, int W, int C,
                                                        int Ho, int Wo, int k, int s, int pad) {
  constexpr int VN = PoolVec<T>::N;
  const int vpp = C / VN;
  const long total = (long)N * Ho * Wo * vpp;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int v = (int)(i % vpp); long r = i / vpp; const int wo = (int)(r % Wo); r /= Wo; const int ho = (int)(r % Ho); const int n = (int)(r / Ho);
    float best[VN]; alignas(8) uint8_t bi[VN];
#pragma unroll
    for (int j = 0; j < VN; ++j) { best[j] = -INFINITY; bi[j] = 0; }
    for (int kh = 0; kh < k; ++kh) for (int kw = 0; kw < k; ++kw) {
      const int h = ho * s - pad + kh, w = wo * s - pad + kw;
      if (h < 0 || h >= H || w < 0 || w >= W) continue;
      float t[VN];
      pool_load(x + (((long)n * H + h) * W + w) * C + v * VN, t);
#pragma unroll
      for (int j = 0; j < VN; ++j) if (t[j] > best[j]) { best[j] = t[j]; bi[j] = (uint8_t)(kh * k + kw); }
    }
    pool_store(y + i * VN, best);
    if (VN == 8) *reinterpret_cast<uint2*>(arg + i * VN) = *reinterpret_cast<const uint2*>(bi);
    else *reinterpret_cast<uint32_t*>(arg + i * VN) = *reinterpret_cast<const uint32_t*>(bi);
  }
}
template <typename T> __global__ void k_maxpool_bwd_vec(const T* __restrict__ dy, const uint8_t* __restrict__ arg, T* __restrict__ dx, int N, int H, int W,
                                                        int C, int Ho, int Wo, int k, int s, int pad) {
  constexpr int VN = PoolVec<T>::N;
  const int vpp = C / VN;
  const long total = (long)N * H * W * vpp;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int v = (int)(i % vpp); long r = i / vpp; const int w = (int)(r % W); r /= W; const int h = (int)(r % H); const int n = (int)(r / H);
    float acc[VN];
#pragma unroll
    for (int j = 0; j < VN; ++j) acc[j] = 0.f;
    const int ho_lo = max(0, (h + pad - k + s) / s), ho_hi = min(Ho - 1, (h + pad) / s);
    const int wo_lo = max(0, (w + pad - k + s) / s), wo_hi = min(Wo - 1, (w + pad) / s);
    for (int ho = ho_lo; ho <= ho_hi; ++ho) for (int wo = wo_lo; wo <= wo_hi; ++wo) {
      const int kh = h + pad - ho * s, kw = w + pad - wo * s;
      if (kh < 0 || kh >= k || kw < 0 || kw >= k) continue;
      const long oi = ((((long)n * Ho + ho) * Wo + wo) * vpp + v) * VN;
      alignas(8) uint8_t a[VN];
      if (VN == 8) *reinterpret_cast<uint2*>(a) = *reinterpret_cast<const uint2*>(arg + oi);
      else *reinterpret_cast<uint32_t*>(a) = *reinterpret_cast<const uint32_t*>(arg + oi);
      float g[VN];
      pool_load(dy + oi, g);
      const int code = kh * k + kw;
#pragma unroll
      for (int j = 0; j < VN; ++j) if (a[j] == code) acc[j] += g[j];
    }
    pool_store(dx + i * VN, acc);
  }
}

// [N, HW, C] -> [N, C] mean; one block per (n, 32-channel group), threads = 32 channels x 8 row lanes
template <typename T> __global__ void k_gap_fwd(const T* __restrict__ x, T* __restrict__ y, int HW, int C) {
  __shared__ float sm[8][33];
  int n = blockIdx.y, c = blockIdx.x * 32 + (threadIdx.x & 31), lane = threadIdx.x >> 5;
  float acc = 0.f;
  if (c < C) for (int r = lane; r < HW; r += 8) acc += to_f(x[((long)n * HW + r) * C + c]);
  sm[lane][threadIdx.x & 31] = acc;
  __syncthreads();
  if (lane == 0 && c < C) {
    float t = 0.f;
    for (int l = 0; l < 8; ++l) t += sm[l][threadIdx.x];
    y[(long)n * C + c] = from_f<T>(t / HW);
  }
}
template <typename T> __global__ void k_gap_bwd(const T* __restrict__ dy, T* __restrict__ dx, long total, int HW, int C) {
  const float inv = 1.f / HW;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    int c = i % C; long n = i / ((long)HW * C);
    dx[i] = from_f<T>(to_f(dy[n * C + c]) * inv);
  }
}

// rows whose argmax equals the label; one warp per row; accumulates into *count (int64, caller zeroes)
template <typename T> __global__ void k_correct_count(const T* __restrict__ x, const long long* __restrict__ target, long long* __restrict__ count, int M, int C) {
  int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  int hit = 0;
  if (row < M) {
    float best = -INFINITY; int bi = 0x7fffffff;
    for (int c = lane; c < C; c += 32) { float v = to_f(x[(long)row * C + c]); if (v > best) { best = v; bi = c; } }
    for (int o = 16; o > 0; o >>= 1) {
      float ob = __shfl_xor_sync(0xffffffffu, best, o); int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (ob > best || (ob == best && oi < bi)) { best = ob; bi = oi; }
    }
    hit = (lane == 0 && bi == (int)target[row]) ? 1 : 0;
  }
  unsigned m = __ballot_sync(0xffffffffu, hit);
  (void)m;
  if (hit) atomicAdd((unsigned long long*)count, 1ull);
}
}  // namespace

// Strided-convolution helpers -----------------------------------------------------------------------------------
namespace {
// up[n, h, w, :] = dy[n, h/2, w/2, :] where h, w are even (and inside dy), 0 elsewhere: the data-gradient of a stride-2
// convolution is the stride-1 data-gradient of this zero-interleaved tensor.  16-byte vectors over the channels.
__global__ void k_upsample2_zero(const uint4* __restrict__ dy, uint4* __restrict__ up, int N, int Ho, int Wo, int Hup, int Wup, int vec_per_pixel) {
  const long total = (long)N * Hup * Wup * vec_per_pixel;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int v = (int)(i % vec_per_pixel); long t = i / vec_per_pixel;
    const int w = (int)(t % Wup); t /= Wup; const int h = (int)(t % Hup); const int n = (int)(t / Hup);
    uint4 val = make_uint4(0, 0, 0, 0);
    if (!(h & 1) && !(w & 1) && (h >> 1) < Ho && (w >> 1) < Wo) val = dy[(((long)n * Ho + (h >> 1)) * Wo + (w >> 1)) * vec_per_pixel + v];
    up[i] = val;
  }
}
// Patch matrix of a convolution over few input channels (the 7x7 / 3-channel ResNet stem): row = output pixel,
// column = (r * S + s) * C + c, zero padded to Kpad columns.  out is [N*Ho*Wo, Kpad]; the stem then is a plain GEMM.
template <typename T>
__global__ void k_im2col_small_c(const T* __restrict__ x, T* __restrict__ out, int N, int H, int W, int C, int R, int S, int stride, int pad,
                                 int Ho, int Wo, int Kpad) {
  // one thread = one 16-byte vector of a patch row (Kpad is a multiple of the K slab, hence of the vector)
  constexpr int VN = 16 / sizeof(T);
  const int vpr = Kpad / VN, Kd = R * S * C;
  const long total = (long)N * Ho * Wo * vpr;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int col0 = (int)(i % vpr) * VN; long t = i / vpr;
    const int wo = (int)(t % Wo); t /= Wo; const int ho = (int)(t % Ho); const int n = (int)(t / Ho);
    alignas(16) T v[VN];
#pragma unroll
    for (int j = 0; j < VN; ++j) {
      const int col = col0 + j;
      T e = from_f<T>(0.f);
      if (col < Kd) {
        const int c = col % C, rs = col / C, s_ = rs % S, r = rs / S;
        const int h = ho * stride - pad + r, w = wo * stride - pad + s_;
        if (h >= 0 && h < H && w >= 0 && w < W) e = x[(((long)n * H + h) * W + w) * C + c];
      }
      v[j] = e;
    }
    *reinterpret_cast<uint4*>(out + i * VN) = *reinterpret_cast<const uint4*>(v);
  }
}
}  // namespace

extern "C" {

// dy: channels-last [N, Ho, Wo, K]; up: [N, Hup, Wup, K] (K * sizeof(element) must be a multiple of 16)
int dcp_upsample2_zero(const void* dy, void* up, int N, int Ho, int Wo, int Hup, int Wup, int K, int dtype, cudaStream_t s) {
  const int es = dtype == DCP_F32 ? 4 : 2;
  if ((K * es) % 16) { dcp_set_error("upsample2_zero: %d channels of %d bytes are not a multiple of 16 bytes", K, es); return -1; }
  const int vpp = K * es / 16;
  const long total = (long)N * Hup * Wup * vpp;
  k_upsample2_zero<<<grid_1d(total, kT), kT, 0, s>>>((const uint4*)dy, (uint4*)up, N, Ho, Wo, Hup, Wup, vpp);
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}
int dcp_im2col_small_c(const void* x, void* out, int N, int H, int W, int C, int R, int S, int stride, int pad, int Ho, int Wo, int Kpad, int dtype,
                       cudaStream_t s) {
  if (Kpad % 8) { dcp_set_error("im2col: Kpad must be a multiple of 8"); return -1; }
  const long total = (long)N * Ho * Wo * Kpad / 4;
  DISPATCH_ACT(dtype, (k_im2col_small_c<T><<<grid_1d(total, kT, 148 * 32), kT, 0, s>>>((const T*)x, (T*)out, N, H, W, C, R, S, stride, pad, Ho, Wo, Kpad)));
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}

}  // extern "C"

extern "C" {

int dcp_relu_fwd(const void* x, void* y, long n, int dtype, cudaStream_t s) {
  DISPATCH_ACT(dtype, (k_relu_fwd<T><<<grid_1d(n, kT), kT, 0, s>>>((const T*)x, (T*)y, n)));
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}
int dcp_relu_bwd(const void* y, const void* dy, void* dx, long n, int dtype, cudaStream_t s) {
  DISPATCH_ACT(dtype, (k_relu_bwd<T><<<grid_1d(n, kT), kT, 0, s>>>((const T*)y, (const T*)dy, (T*)dx, n)));
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}
int dcp_add_relu_fwd(const void* a, const void* b, void* y, long n, int dtype, cudaStream_t s) {
  DISPATCH_ACT(dtype, (k_add_relu_fwd<T><<<grid_1d(n, kT), kT, 0, s>>>((const T*)a, (const T*)b, (T*)y, n)));
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}
int dcp_dropout(const void* x, void* y, long n, float p, unsigned long long seed, unsigned long long stream_id, const void* step, int mode, long inner,
                long hw, int dtype, cudaStream_t s) {
  DISPATCH_ACT(dtype, (k_dropout<T><<<grid_1d(n, kT), kT, 0, s>>>((const T*)x, (T*)y, n, p, seed, stream_id, (const long long*)step, mode, inner, hw)));
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}
int dcp_maxpool_fwd(const void* x, void* y, void* arg, int N, int H, int W, int C, int Ho, int Wo, int k, int st, int pad, int out_nchw, int dtype, cudaStream_t s) {
  long total = (long)N * Ho * Wo * C;
  const int vn = dtype == DCP_F32 ? 4 : 8;
  if (!out_nchw && C % vn == 0 && ((uintptr_t)x % 16 == 0) && ((uintptr_t)y % 16 == 0) && ((uintptr_t)arg % 8 == 0)) {
    DISPATCH_ACT(dtype, (k_maxpool_fwd_vec<T><<<grid_1d(total / vn, kT), kT, 0, s>>>((const T*)x, (T*)y, (uint8_t*)arg, N, H, W, C, Ho, Wo, k, st, pad)));
    DCP_CUDA_CHECK(cudaGetLastError()); return 0;
  }
  DISPATCH_ACT(dtype, (k_maxpool_fwd<T><<<grid_1d(total, kT), kT, 0, s>>>((const T*)x, (T*)y, (uint8_t*)arg, N, H, W, C, Ho, Wo, k, st, pad, out_nchw)));
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}
int dcp_maxpool_bwd(const void* dy, const void* arg, void* dx, int N, int H, int W, int C, int Ho, int Wo, int k, int st, int pad, int out_nchw, int dtype, cudaStream_t s) {
  long total = (long)N * H * W * C;
  const int vn = dtype == DCP_F32 ? 4 : 8;
  if (!out_nchw && C % vn == 0 && ((uintptr_t)dy % 16 == 0) && ((uintptr_t)dx % 16 == 0) && ((uintptr_t)arg % 8 == 0)) {
    DISPATCH_ACT(dtype, (k_maxpool_bwd_vec<T><<<grid_1d(total / vn, kT), kT, 0, s>>>((const T*)dy, (const uint8_t*)arg, (T*)dx, N, H, W, C, Ho, Wo, k, st, pad)));
    DCP_CUDA_CHECK(cudaGetLastError()); return 0;
  }
  DISPATCH_ACT(dtype, (k_maxpool_bwd<T><<<grid_1d(total, kT), kT, 0, s>>>((const T*)dy, (const uint8_t*)arg, (T*)dx, N, H, W, C, Ho, Wo, k, st, pad, out_nchw)));
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}
int dcp_gap_fwd(const void* x, void* y, int N, int HW, int C, int dtype, cudaStream_t s) {
  dim3 grid(ceil_div(C, 32), N);
  DISPATCH_ACT(dtype, (k_gap_fwd<T><<<grid, 256, 0, s>>>((const T*)x, (T*)y, HW, C)));
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}
int dcp_gap_bwd(const void* dy, void* dx, int N, int HW, int C, int dtype, cudaStream_t s) {
  long total = (long)N * HW * C;
  DISPATCH_ACT(dtype, (k_gap_bwd<T><<<grid_1d(total, kT), kT, 0, s>>>((const T*)dy, (T*)dx, total, HW, C)));
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}
int dcp_correct_count(const void* x, const void* target, void* count, int M, int C, int dtype, cudaStream_t s) {
  DCP_CUDA_CHECK(cudaMemsetAsync(count, 0, 8, s));
  DISPATCH_ACT(dtype, (k_correct_count<T><<<ceil_div(M, 8), 256, 0, s>>>((const T*)x, (const long long*)target, (long long*)count, M, C)));
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}

}  // extern "C"
