// Fused classification head kernels: log_softmax, nll_loss, and the fused cross-entropy
// (log_softmax + nll forward; (softmax - onehot) * g backward) - call sites K13/K14 of the reference
// (F.log_softmax at ref:main.py:44, F.nll_loss mean at :61 and sum at :81-82).  One CTA per row; rows
// may be as short as 10 classes (ConvNet) or as long as a 50k vocabulary (GPT-2).
#include "ops/ops_common.cuh"

namespace {
constexpr int kT = 128;

template <typename T> __device__ __forceinline__ float row_lse(const T* __restrict__ x, int C, float* red) {
  float m = -INFINITY;
  for (int c = threadIdx.x; c < C; c += blockDim.x) m = fmaxf(m, to_f(x[c]));
  m = block_max(m, red);
  float s = 0.f;
  for (int c = threadIdx.x; c < C; c += blockDim.x) s += __expf(to_f(x[c]) - m);
  s = block_sum(s, red);
  return m + __logf(s);
}

template <typename T> __global__ void k_log_softmax_fwd(const T* __restrict__ x, T* __restrict__ y, int C) {
  __shared__ float red[32];
  const T* xr = x + (long)blockIdx.x * C; T* yr = y + (long)blockIdx.x * C;
  float lse = row_lse(xr, C, red);
  for (int c = threadIdx.x; c < C; c += blockDim.x) yr[c] = from_f<T>(to_f(xr[c]) - lse);
}
// dx = dy - exp(y) * sum(dy)
template <typename T> __global__ void k_log_softmax_bwd(const T* __restrict__ y, const T* __restrict__ dy, T* __restrict__ dx, int C) {
  __shared__ float red[32];
  const long o = (long)blockIdx.x * C;
  float s = 0.f;
  for (int c = threadIdx.x; c < C; c += blockDim.x) s += to_f(dy[o + c]);
  s = block_sum(s, red);
  for (int c = threadIdx.x; c < C; c += blockDim.x) dx[o + c] = from_f<T>(to_f(dy[o + c]) - __expf(to_f(y[o + c])) * s);
}

// loss += -scale * logp[row, target[row]]   (scale = 1/M for mean, 1 for sum); *loss zeroed by the launcher
template <typename T> __global__ void k_nll_fwd(const T* __restrict__ logp, const long long* __restrict__ target, float* __restrict__ loss, int M, int C, float scale) {
  __shared__ float red[32];
  float a = 0.f;
  for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < M; r += gridDim.x * blockDim.x) {
    long long t = target[r];
    if (t >= 0 && t < C) a -= to_f(logp[(long)r * C + t]);
  }
  a = block_sum(a, red);
  if (threadIdx.x == 0) atomicAdd(loss, a * scale);
}
// dlogp[row, c] = -(c == target) * g * scale
template <typename T> __global__ void k_nll_bwd(const long long* __restrict__ target, const float* __restrict__ gout, T* __restrict__ dlogp, int M, int C, float scale) {
  const float g = -(*gout) * scale;
  const long total = (long)M * C;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    int c = i % C; long r = i / C;
    dlogp[i] = from_f<T>(target[r] == c ? g : 0.f);
  }
}

// fused: per-row lse saved for backward; loss accumulated with one atomic per CTA
template <typename T> __global__ void k_ce_fwd(const T* __restrict__ x, const long long* __restrict__ target, float* __restrict__ loss,
                                               float* __restrict__ lse_out, int C, float scale) {
  __shared__ float red[32];
  const T* xr = x + (long)blockIdx.x * C;
  float lse = row_lse(xr, C, red);
  if (threadIdx.x == 0) {
    lse_out[blockIdx.x] = lse;
    long long t = target[blockIdx.x];
    if (t >= 0 && t < C) atomicAdd(loss, (lse - to_f(xr[t])) * scale);
  }
}
// dx = (softmax - onehot) * g * scale; dx may alias x
template <typename T> __global__ void k_ce_bwd(const T* x, const long long* __restrict__ target, const float* __restrict__ lse,
                                               const float* __restrict__ gout, T* dx, int C, float scale) {
  const long o = (long)blockIdx.x * C;
  const float g = (*gout) * scale, l = lse[blockIdx.x];
  const long long t = target[blockIdx.x];
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float p = __expf(to_f(x[o + c]) - l);
    dx[o + c] = from_f<T>((p - (c == t ? 1.f : 0.f)) * g);
  }
}

// ---- long rows (a 50 k vocabulary): 16-byte loads, ONE pass with an online (max, sum) pair per thread -------------------
template <typename T> struct CeVec;
template <> struct CeVec<float> { static constexpr int N = 4; __device__ static void load(const float* p, float* o) { const float4 v = *reinterpret_cast<const float4*>(p); o[0] = v.x; o[1] = v.y; o[2] = v.z; o[3] = v.w; }
                                  __device__ static void store(float* p, const float* o) { *reinterpret_cast<float4*>(p) = make_float4(o[0], o[1], o[2], o[3]); } };
template <> struct CeVec<__nv_bfloat16> {
  static constexpr int N = 8;
  __device__ static void load(const __nv_bfloat16* p, float* o) {
    const uint4 v = *reinterpret_cast<const uint4*>(p);
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
    for (int i = 0; i < 4; ++i) { o[2 * i] = __low2float(h[i]); o[2 * i + 1] = __high2float(h[i]); }
  }
  __device__ static void store(__nv_bfloat16* p, const float* o) {
    uint4 v; __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&v);
#pragma unroll
    for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(o[2 * i], o[2 * i + 1]);
    *reinterpret_cast<uint4*>(p) = v;
  }
};
__device__ __forceinline__ void lse_merge(float& m, float& s, float m2, float s2) {
  const float M = fmaxf(m, m2);
  s = M == -INFINITY ? 0.f : s * __expf(m - M) + s2 * __expf(m2 - M);
  m = M;
}
template <typename T> __global__ void __launch_bounds__(1024) k_ce_fwd_vec(const T* __restrict__ x, const long long* __restrict__ target, float* __restrict__ loss,
                                                                           float* __restrict__ lse_out, int C, float scale) {
  constexpr int VN = CeVec<T>::N;
  __shared__ float rm[32], rs[32];
  const T* xr = x + (long)blockIdx.x * C;
  const int nvec = C / VN;
  float m = -INFINITY, s = 0.f;
  for (int vi = threadIdx.x; vi < nvec; vi += blockDim.x) {
    float v[VN];
    CeVec<T>::load(xr + vi * VN, v);
    float mv = v[0];
#pragma unroll
    for (int i = 1; i < VN; ++i) mv = fmaxf(mv, v[i]);
    if (mv > m) { s *= __expf(m - mv); m = mv; }
#pragma unroll
    for (int i = 0; i < VN; ++i) s += __expf(v[i] - m);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) lse_merge(m, s, __shfl_xor_sync(0xffffffffu, m, o), __shfl_xor_sync(0xffffffffu, s, o));
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
  if (lane == 0) { rm[w] = m; rs[w] = s; }
  __syncthreads();
  if (w == 0) {
    m = lane < nw ? rm[lane] : -INFINITY; s = lane < nw ? rs[lane] : 0.f;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) lse_merge(m, s, __shfl_xor_sync(0xffffffffu, m, o), __shfl_xor_sync(0xffffffffu, s, o));
    if (lane == 0) {
      const float lse = m + __logf(s);
      lse_out[blockIdx.x] = lse;
      const long long t = target[blockIdx.x];
      if (t >= 0 && t < C) atomicAdd(loss, (lse - to_f(xr[t])) * scale);
    }
  }
}
template <typename T> __global__ void __launch_bounds__(1024) k_ce_bwd_vec(const T* x, const long long* __restrict__ target, const float* __restrict__ lse,
                                                                           const float* __restrict__ gout, T* dx, int C, float scale) {
  constexpr int VN = CeVec<T>::N;
  const long o = (long)blockIdx.x * C;
  const float g = (*gout) * scale, l = lse[blockIdx.x];
  const int t = (int)target[blockIdx.x], nvec = C / VN;
  for (int vi = threadIdx.x; vi < nvec; vi += blockDim.x) {
    float v[VN];
    CeVec<T>::load(x + o + vi * VN, v);
#pragma unroll
    for (int i = 0; i < VN; ++i) v[i] = (__expf(v[i] - l) - (vi * VN + i == t ? 1.f : 0.f)) * g;
    CeVec<T>::store(dx + o + vi * VN, v);
  }
}
template <typename T> bool ce_vec_ok(const void* x, const void* dx, int C) {
  return C >= 2048 && C % CeVec<T>::N == 0 && (uintptr_t)x % 16 == 0 && (dx == nullptr || (uintptr_t)dx % 16 == 0);
}
}  // namespace

extern "C" {
int dcp_log_softmax_fwd(const void* x, void* y, int M, int C, int dtype, cudaStream_t s) {
  DISPATCH_ACT(dtype, (k_log_softmax_fwd<T><<<M, kT, 0, s>>>((const T*)x, (T*)y, C)));
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}
int dcp_log_softmax_bwd(const void* y, const void* dy, void* dx, int M, int C, int dtype, cudaStream_t s) {
  DISPATCH_ACT(dtype, (k_log_softmax_bwd<T><<<M, kT, 0, s>>>((const T*)y, (const T*)dy, (T*)dx, C)));
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}
int dcp_nll_fwd(const void* logp, const void* target, float* loss, int M, int C, float scale, int dtype, cudaStream_t s) {
  DCP_CUDA_CHECK(cudaMemsetAsync(loss, 0, sizeof(float), s));
  DISPATCH_ACT(dtype, (k_nll_fwd<T><<<grid_1d(M, 256, 64), 256, 0, s>>>((const T*)logp, (const long long*)target, loss, M, C, scale)));
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}
int dcp_nll_bwd(const void* target, const float* gout, void* dlogp, int M, int C, float scale, int dtype, cudaStream_t s) {
  DISPATCH_ACT(dtype, (k_nll_bwd<T><<<grid_1d((long)M * C, 256), 256, 0, s>>>((const long long*)target, gout, (T*)dlogp, M, C, scale)));
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}
int dcp_ce_fwd(const void* x, const void* target, float* loss, float* lse, int M, int C, float scale, int dtype, cudaStream_t s) {
  DCP_CUDA_CHECK(cudaMemsetAsync(loss, 0, sizeof(float), s));
  DISPATCH_ACT(dtype, (ce_vec_ok<T>(x, nullptr, C) ? k_ce_fwd_vec<T><<<M, 1024, 0, s>>>((const T*)x, (const long long*)target, loss, lse, C, scale)
                                                    : k_ce_fwd<T><<<M, C >= 1024 ? 512 : kT, 0, s>>>((const T*)x, (const long long*)target, loss, lse, C, scale)));
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}
int dcp_ce_bwd(const void* x, const void* target, const float* lse, const float* gout, void* dx, int M, int C, float scale, int dtype, cudaStream_t s) {
  DISPATCH_ACT(dtype, (ce_vec_ok<T>(x, dx, C) ? k_ce_bwd_vec<T><<<M, 1024, 0, s>>>((const T*)x, (const long long*)target, lse, gout, (T*)dx, C, scale)
                                               : k_ce_bwd<T><<<M, C >= 1024 ? 512 : kT, 0, s>>>((const T*)x, (const long long*)target, lse, gout, (T*)dx, C, scale)));
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}
}  // extern "C"
