// CUDA-core convolution / linear kernels (forward, data-gradient, weight-gradient) on the generic
// implicit-GEMM skeleton.  They serve the shapes where tensor cores do not pay - conv1 of the reference
// model (C_in = 1, K = 9; ref:main.py:32), fc2 (N = 10; ref:main.py:43), the 3-channel ResNet stem - and
// act as the numerics cross-check for the tcgen05 path (csrc/ops/gemm_tc.cu, conv_tc.cu).
// Layouts: activations channels-last [N,H,W,C]; conv weights [K,R,S,C]; linear weights [N_out, K_in].
#include "ops/simt_gemm.cuh"

struct ConvP { int N, H, W, C, K, R, S, stride, pad, Ho, Wo; };

namespace {

template <typename T> struct ConvFwdA {
  const T* x; ConvP p;
  __device__ float operator()(int m, int kk) const {
    int c = kk % p.C; int rs = kk / p.C; int s = rs % p.S, r = rs / p.S;
    int wo = m % p.Wo; int t = m / p.Wo; int ho = t % p.Ho, n = t / p.Ho;
    int h = ho * p.stride - p.pad + r, w = wo * p.stride - p.pad + s;
    if (h < 0 || h >= p.H || w < 0 || w >= p.W) return 0.f;
    return to_f(x[(((long)n * p.H + h) * p.W + w) * p.C + c]);
  }
};
template <typename T> struct ConvFwdB {  // B[kk][k] = w[k][kk]
  const T* w; int RSC;
  __device__ float operator()(int kk, int k) const { return to_f(w[(long)k * RSC + kk]); }
};
template <typename T> struct StoreBiasAct {
  T* y; const T* bias; int ld; int relu;
  __device__ void operator()(int m, int n, float acc) const {
    if (bias) acc += to_f(bias[n]);
    if (relu) acc = fmaxf(acc, 0.f);
    y[(long)m * ld + n] = from_f<T>(acc);
  }
};
struct StoreAtomicF32 {
  float* y; int ld;
  __device__ void operator()(int m, int n, float acc) const { atomicAdd(&y[(long)m * ld + n], acc); }
};

// dgrad: dx[n,h,w,c] = sum_{r,s,k} dy[n,(h+pad-r)/st,(w+pad-s)/st,k] * w[k,r,s,c]
template <typename T> struct ConvDgradA {
  const T* dy; ConvP p;
  __device__ float operator()(int m, int kk) const {
    int k = kk % p.K; int rs = kk / p.K; int s = rs % p.S, r = rs / p.S;
    int w = m % p.W; int t = m / p.W; int h = t % p.H, n = t / p.H;
    int hn = h + p.pad - r, wn = w + p.pad - s;
    if (hn < 0 || wn < 0 || hn % p.stride || wn % p.stride) return 0.f;
    int ho = hn / p.stride, wo = wn / p.stride;
    if (ho >= p.Ho || wo >= p.Wo) return 0.f;
    return to_f(dy[(((long)n * p.Ho + ho) * p.Wo + wo) * p.K + k]);
  }
};
template <typename T> struct ConvDgradB {  // B[(r,s,k)][c] = w[k][r][s][c]
  const T* w; ConvP p;
  __device__ float operator()(int kk, int c) const {
    int k = kk % p.K; int rs = kk / p.K;
    return to_f(w[((long)k * p.R * p.S + rs) * p.C + c]);
  }
};
// wgrad: dw[k][(r,s,c)] = sum_pixels dy[pixel][k] * x[pixel shifted][c]
template <typename T> struct ConvWgradA {
  const T* dy; int K;
  __device__ float operator()(int k, int pix) const { return to_f(dy[(long)pix * K + k]); }
};
template <typename T> struct ConvWgradB {
  const T* x; ConvP p;
  __device__ float operator()(int pix, int j) const {
    int c = j % p.C; int rs = j / p.C; int s = rs % p.S, r = rs / p.S;
    int wo = pix % p.Wo; int t = pix / p.Wo; int ho = t % p.Ho, n = t / p.Ho;
    int h = ho * p.stride - p.pad + r, w = wo * p.stride - p.pad + s;
    if (h < 0 || h >= p.H || w < 0 || w >= p.W) return 0.f;
    return to_f(x[(((long)n * p.H + h) * p.W + w) * p.C + c]);
  }
};

// plain row-major loaders: element (i, j) at base[i*si + j*sj]
template <typename T> struct Strided {
  const T* base; long si, sj;
  __device__ float operator()(int i, int j) const { return to_f(base[i * si + j * sj]); }
};

// column sums of a [M, C] matrix into float out[C] (+=), used for bias gradients
template <typename T> __global__ void k_colsum(const T* __restrict__ x, float* __restrict__ out, long M, int C, long rows_per_cta) {
  __shared__ float sm[8][33];
  const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5, c = blockIdx.x * 32 + cx;
  const long r0 = (long)blockIdx.y * rows_per_cta, r1 = min(M, r0 + rows_per_cta);
  float a = 0.f;
  if (c < C) for (long r = r0 + ry; r < r1; r += 8) a += to_f(x[r * C + c]);
  sm[ry][cx] = a;
  __syncthreads();
  if (ry == 0 && c < C) {
    for (int i = 1; i < 8; ++i) a += sm[i][cx];
    atomicAdd(&out[c], a);
  }
}
// 16-byte variant (C a multiple of the vector, aligned base): a thread owns 8 bf16 / 4 f32 channels, 4 rows in flight; the row lanes
// that share a warp fold by shuffle, one shared-memory atomic per channel and warp, one global atomic per channel and CTA.
template <typename T> __global__ void __launch_bounds__(256) k_colsum_vec(const T* __restrict__ x, float* __restrict__ out, long M, int C, int cl,
                                                                          long rows_per_cta) {
  constexpr int VN = 16 / sizeof(T);
  __shared__ float sacc[32 * 8];
  const int rl = 256 / cl, lx = threadIdx.x % cl, ly = threadIdx.x / cl;
  const int vcol = blockIdx.x * cl + lx;
  const bool live = vcol * VN < C;
  for (int i = threadIdx.x; i < cl * VN; i += 256) sacc[i] = 0.f;
  __syncthreads();
  float a[VN];
#pragma unroll
  for (int i = 0; i < VN; ++i) a[i] = 0.f;
  const long r0 = (long)blockIdx.y * rows_per_cta, r1 = min(M, r0 + rows_per_cta);
  if (live) {
    for (long r = r0 + ly; r < r1; r += 4L * rl) {
      alignas(16) T v[4][VN];
      float wgt[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const long rr = r + (long)u * rl;
        wgt[u] = rr < r1 ? 1.f : 0.f;
        *reinterpret_cast<uint4*>(v[u]) = *reinterpret_cast<const uint4*>(x + min(rr, r1 - 1) * C + (long)vcol * VN);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u)
#pragma unroll
        for (int i = 0; i < VN; ++i) a[i] += wgt[u] * to_f(v[u][i]);
    }
  }
  const bool fold = cl < 32 && (32 % cl) == 0;
  if (fold) {
#pragma unroll
    for (int i = 0; i < VN; ++i)
      for (int o = cl; o < 32; o <<= 1) a[i] += __shfl_xor_sync(0xffffffffu, a[i], o);
  }
  if (live && (!fold || (threadIdx.x & 31) < cl)) {
#pragma unroll
    for (int i = 0; i < VN; ++i) atomicAdd(&sacc[lx * VN + i], a[i]);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < cl * VN; i += 256) {
    const int c = blockIdx.x * cl * VN + i;
    if (c < C) atomicAdd(&out[c], sacc[i]);
  }
}
// launches the vector variant when the shape allows; returns false otherwise
template <typename T> bool colsum_vec(const T* x, float* out, long M, int C, cudaStream_t s) {
  constexpr int VN = 16 / sizeof(T);
  if (C % VN || (reinterpret_cast<uintptr_t>(x) & 15) || M * (long)C < 65536) return false;
  const int ncl = C / VN;
  int cl = ncl < 32 ? ncl : 32;
  while (256 % cl) --cl;
  const int gx = (ncl + cl - 1) / cl, rl = 256 / cl;
  long want = (148L * 2 + gx - 1) / gx, maxy = (M + rl * 4L - 1) / (rl * 4L);
  long gy = want < maxy ? want : maxy; if (gy < 1) gy = 1; if (gy > 65535) gy = 65535;
  k_colsum_vec<T><<<dim3(gx, (unsigned)gy), 256, 0, s>>>(x, out, M, C, cl, (M + gy - 1) / gy);
  return true;
}
__global__ void __launch_bounds__(256) k_cast_bf16_vec(const float* __restrict__ in, __nv_bfloat16* __restrict__ out, long n8) {
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (long)gridDim.x * blockDim.x) {
    const float4 a = reinterpret_cast<const float4*>(in)[2 * i], b = reinterpret_cast<const float4*>(in)[2 * i + 1];
    const __nv_bfloat162 p0 = __floats2bfloat162_rn(a.x, a.y), p1 = __floats2bfloat162_rn(a.z, a.w), p2 = __floats2bfloat162_rn(b.x, b.y), p3 = __floats2bfloat162_rn(b.z, b.w);
    reinterpret_cast<uint4*>(out)[i] = make_uint4(*reinterpret_cast<const uint32_t*>(&p0), *reinterpret_cast<const uint32_t*>(&p1),
                                                  *reinterpret_cast<const uint32_t*>(&p2), *reinterpret_cast<const uint32_t*>(&p3));
  }
}
template <typename T> __global__ void k_cast_from_f32(const float* __restrict__ in, T* __restrict__ out, long n) {
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) out[i] = from_f<T>(in[i]);
}

}  // namespace

extern "C" {

int dcp_conv2d_fwd_simt(const void* x, const void* w, const void* bias, void* y, ConvP p, int relu, int dtype, cudaStream_t s) {
  const int M = p.N * p.Ho * p.Wo, Kd = p.R * p.S * p.C;
  DISPATCH_ACT(dtype, (launch_simt_gemm(ConvFwdA<T>{(const T*)x, p}, ConvFwdB<T>{(const T*)w, Kd}, StoreBiasAct<T>{(T*)y, (const T*)bias, p.K, relu}, M, p.K, Kd, 1, s)));
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}
int dcp_conv2d_dgrad_simt(const void* dy, const void* w, void* dx, ConvP p, int dtype, cudaStream_t s) {
  const int M = p.N * p.H * p.W, Kd = p.R * p.S * p.K;
  DISPATCH_ACT(dtype, (launch_simt_gemm(ConvDgradA<T>{(const T*)dy, p}, ConvDgradB<T>{(const T*)w, p}, StoreBiasAct<T>{(T*)dx, nullptr, p.C, 0}, M, p.C, Kd, 1, s)));
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}
// dw32: float [K, R*S*C] accumulator (zeroed here).  db32: float [K] or null.
int dcp_conv2d_wgrad_simt(const void* dy, const void* x, float* dw32, float* db32, ConvP p, int dtype, cudaStream_t s) {
  const int Mo = p.K, No = p.R * p.S * p.C, Kd = p.N * p.Ho * p.Wo;
  DCP_CUDA_CHECK(cudaMemsetAsync(dw32, 0, sizeof(float) * (size_t)Mo * No, s));
  DISPATCH_ACT(dtype, (launch_simt_gemm(ConvWgradA<T>{(const T*)dy, p.K}, ConvWgradB<T>{(const T*)x, p}, StoreAtomicF32{dw32, No}, Mo, No, Kd, simt_ksplit(Mo, No, Kd), s)));
  if (db32) {
    DCP_CUDA_CHECK(cudaMemsetAsync(db32, 0, sizeof(float) * p.K, s));
    long gy = Kd / 256 + 1; if (gy > 592) gy = 592;
    long rpc = (Kd + gy - 1) / gy;
    DISPATCH_ACT(dtype, (k_colsum<T><<<dim3(ceil_div(p.K, 32), (unsigned)gy), 256, 0, s>>>((const T*)dy, db32, Kd, p.K, rpc)));
  }
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}

// y[M,N] = x[M,K] . w[N,K]^T + b
int dcp_linear_fwd_simt(const void* x, const void* w, const void* bias, void* y, int M, int N, int K, int relu, int dtype, cudaStream_t s) {
  DISPATCH_ACT(dtype, (launch_simt_gemm(Strided<T>{(const T*)x, K, 1}, Strided<T>{(const T*)w, 1, K}, StoreBiasAct<T>{(T*)y, (const T*)bias, N, relu}, M, N, K, 1, s)));
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}
// dx[M,K] = dy[M,N] . w[N,K]
int dcp_linear_dgrad_simt(const void* dy, const void* w, void* dx, int M, int N, int K, int dtype, cudaStream_t s) {
  DISPATCH_ACT(dtype, (launch_simt_gemm(Strided<T>{(const T*)dy, N, 1}, Strided<T>{(const T*)w, K, 1}, StoreBiasAct<T>{(T*)dx, nullptr, K, 0}, M, K, N, 1, s)));
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}
// dw32[N,K] = dy[M,N]^T . x[M,K] (float accumulator, zeroed here); db32[N] = column sums of dy
int dcp_linear_wgrad_simt(const void* dy, const void* x, float* dw32, float* db32, int M, int N, int K, int dtype, cudaStream_t s) {
  DCP_CUDA_CHECK(cudaMemsetAsync(dw32, 0, sizeof(float) * (size_t)N * K, s));
  DISPATCH_ACT(dtype, (launch_simt_gemm(Strided<T>{(const T*)dy, 1, N}, Strided<T>{(const T*)x, K, 1}, StoreAtomicF32{dw32, K}, N, K, M, simt_ksplit(N, K, M), s)));
  if (db32) {
    DCP_CUDA_CHECK(cudaMemsetAsync(db32, 0, sizeof(float) * N, s));
    long gy = M / 256 + 1; if (gy > 592) gy = 592;
    long rpc = (M + gy - 1) / gy;
    DISPATCH_ACT(dtype, (k_colsum<T><<<dim3(ceil_div(N, 32), (unsigned)gy), 256, 0, s>>>((const T*)dy, db32, M, N, rpc)));
  }
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}
// float column sums of [M, C] (+= into zeroed out)
int dcp_colsum(const void* x, float* out, long M, int C, int dtype, cudaStream_t s) {
  DCP_CUDA_CHECK(cudaMemsetAsync(out, 0, sizeof(float) * C, s));
  bool done = false;
  DISPATCH_ACT(dtype, (done = colsum_vec<T>((const T*)x, out, M, C, s)));
  if (done) { DCP_CUDA_CHECK(cudaGetLastError()); return 0; }
  long gy = M / 256 + 1; if (gy > 592) gy = 592;
  long rpc = (M + gy - 1) / gy;
  DISPATCH_ACT(dtype, (k_colsum<T><<<dim3(ceil_div(C, 32), (unsigned)gy), 256, 0, s>>>((const T*)x, out, M, C, rpc)));
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}
int dcp_cast_from_f32(const float* in, void* out, long n, int dtype, cudaStream_t s) {
  if (dtype == DCP_BF16 && n >= 8 && ((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(out)) & 15) == 0) {
    // eight elements per thread: two 16-byte loads, one 16-byte store; the ragged tail goes through the scalar kernel
    const long n8 = n / 8;
    k_cast_bf16_vec<<<grid_1d(n8, 256), 256, 0, s>>>(in, static_cast<__nv_bfloat16*>(out), n8);
    if (n % 8) k_cast_from_f32<__nv_bfloat16><<<1, 32, 0, s>>>(in + n8 * 8, static_cast<__nv_bfloat16*>(out) + n8 * 8, n % 8);
    DCP_CUDA_CHECK(cudaGetLastError()); return 0;
  }
  DISPATCH_ACT(dtype, (k_cast_from_f32<T><<<grid_1d(n, 256), 256, 0, s>>>(in, (T*)out, n)));
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}

}  // extern "C"
