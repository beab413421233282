// Transformer-block kernels for the GPT-2 north-star config (BASELINE.json configs[4]; not in the reference):
// LayerNorm forward/backward (row statistics in fp32, one CTA per row) and tanh-GELU forward/backward.
// Activations f32 or bf16; LayerNorm parameters fp32.
#include "ops/ops_common.cuh"

namespace {
constexpr int kT = 256;

template <typename T>
__global__ void __launch_bounds__(kT) k_layernorm_fwd(const T* __restrict__ x, T* __restrict__ y, const float* __restrict__ g, const float* __restrict__ b,
                                                      float* __restrict__ mean_out, float* __restrict__ rstd_out, int C, float eps) {
  __shared__ float red[32];
  const long o = (long)blockIdx.x * C;
  float s = 0.f;
  for (int c = threadIdx.x; c < C; c += blockDim.x) s += to_f(x[o + c]);
  const float mean = block_sum(s, red) / C;
  float v = 0.f;
  for (int c = threadIdx.x; c < C; c += blockDim.x) { const float d = to_f(x[o + c]) - mean; v += d * d; }
  const float rstd = rsqrtf(block_sum(v, red) / C + eps);
  if (threadIdx.x == 0) { mean_out[blockIdx.x] = mean; rstd_out[blockIdx.x] = rstd; }
  for (int c = threadIdx.x; c < C; c += blockDim.x) y[o + c] = from_f<T>((to_f(x[o + c]) - mean) * rstd * g[c] + b[c]);
}

// dx = rstd * (gy - mean(gy) - xhat * mean(gy * xhat)) with gy = dy * gamma; dgamma/dbeta accumulate (caller zeroes)
template <typename T>
__global__ void __launch_bounds__(kT) k_layernorm_bwd(const T* __restrict__ x, const T* __restrict__ dy, const float* __restrict__ g,
                                                      const float* __restrict__ mean, const float* __restrict__ rstd, T* __restrict__ dx,
                                                      float* __restrict__ dgamma, float* __restrict__ dbeta, int C, int rows, int rows_per_cta) {
  __shared__ float red[32];
  extern __shared__ float acc[];   // [2][C] partial dgamma / dbeta of this CTA
  for (int c = threadIdx.x; c < 2 * C; c += blockDim.x) acc[c] = 0.f;
  __syncthreads();
  const int r0 = blockIdx.x * rows_per_cta, r1 = min(rows, r0 + rows_per_cta);
  for (int r = r0; r < r1; ++r) {
    const long o = (long)r * C;
    const float mu = mean[r], rs = rstd[r];
    float s1 = 0.f, s2 = 0.f;
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
      const float gy = to_f(dy[o + c]) * g[c], xh = (to_f(x[o + c]) - mu) * rs;
      s1 += gy; s2 += gy * xh;
    }
    s1 = block_sum(s1, red) / C; s2 = block_sum(s2, red) / C;
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
      const float d = to_f(dy[o + c]), xh = (to_f(x[o + c]) - mu) * rs;
      dx[o + c] = from_f<T>(rs * (d * g[c] - s1 - xh * s2));
      acc[c] += d * xh; acc[C + c] += d;       // thread-private columns: no race
    }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += blockDim.x) { atomicAdd(&dgamma[c], acc[c]); atomicAdd(&dbeta[c], acc[C + c]); }
}

// ---- warp-per-row variants (C a multiple of the 16-byte vector, C <= 32 * kVPL vectors): one warp owns a row, reads it ONCE
// with 16-byte loads, keeps it in registers, reduces with shuffles (no block barriers); a CTA of 8 warps walks rows with a
// grid stride.  Backward: each lane owns fixed columns, so the dgamma / dbeta partial sums stay in registers over all rows
// of the warp, meet in shared memory once per CTA and leave with one atomic per column and CTA.
template <typename T> struct Vec16;
template <> struct Vec16<float> { static constexpr int N = 4; __device__ static void load(const float* p, float* o) { const float4 v = *reinterpret_cast<const float4*>(p); o[0] = v.x; o[1] = v.y; o[2] = v.z; o[3] = v.w; }
                                  __device__ static void store(float* p, const float* o) { *reinterpret_cast<float4*>(p) = make_float4(o[0], o[1], o[2], o[3]); } };
template <> struct Vec16<__nv_bfloat16> {
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
constexpr int kMaxVPL = 4;   // 16-byte vectors per lane (template parameter VPL = ceil(C / vector / 32)): C <= 1024 (bf16) / 512 (f32)

// res / sum_out (both or neither): the row normalised is x + res, and that sum is also written out - the residual add of a
// pre-LN transformer block fused into the LayerNorm that follows it.
template <typename T, int kVPL>
__global__ void __launch_bounds__(kT) k_layernorm_fwd_warp(const T* __restrict__ x, T* __restrict__ y, const float* __restrict__ g, const float* __restrict__ b,
                                                           float* __restrict__ mean_out, float* __restrict__ rstd_out, int C, int rows, float eps,
                                                           const T* __restrict__ res = nullptr, T* __restrict__ sum_out = nullptr) {
  constexpr int VN = Vec16<T>::N;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nvec = C / VN;
  const float invC = 1.f / (float)C;
  for (int r = blockIdx.x * (kT / 32) + warp; r < rows; r += gridDim.x * (kT / 32)) {
    const T* xr = x + (long)r * C;
    float v[kVPL][VN];
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < kVPL; ++k) {
      const int vi = lane + 32 * k;
      if (vi < nvec) {
        Vec16<T>::load(xr + vi * VN, v[k]);
        if (res != nullptr) {
          float rr[VN];
          Vec16<T>::load(res + (long)r * C + vi * VN, rr);
#pragma unroll
          for (int i = 0; i < VN; ++i) v[k][i] += rr[i];
          Vec16<T>::store(sum_out + (long)r * C + vi * VN, v[k]);
#pragma unroll
          for (int i = 0; i < VN; ++i) v[k][i] = to_f(from_f<T>(v[k][i]));   // normalise what later kernels will read (the rounded sum)
        }
#pragma unroll
        for (int i = 0; i < VN; ++i) s += v[k][i];
      }
    }
    const float mean = warp_sum(s) * invC;
    float q = 0.f;
#pragma unroll
    for (int k = 0; k < kVPL; ++k)
      if (lane + 32 * k < nvec) {
#pragma unroll
        for (int i = 0; i < VN; ++i) { const float d = v[k][i] - mean; q += d * d; }
      }
    const float rstd = rsqrtf(warp_sum(q) * invC + eps);
    if (lane == 0) { mean_out[r] = mean; rstd_out[r] = rstd; }
#pragma unroll
    for (int k = 0; k < kVPL; ++k) {
      const int vi = lane + 32 * k;
      if (vi < nvec) {
        float o[VN];
#pragma unroll
        for (int i = 0; i < VN; ++i) o[i] = (v[k][i] - mean) * rstd * g[vi * VN + i] + b[vi * VN + i];
        Vec16<T>::store(y + (long)r * C + vi * VN, o);
      }
    }
  }
}

template <typename T, int kVPL>
__global__ void __launch_bounds__(kT) k_layernorm_bwd_warp(const T* __restrict__ x, const T* __restrict__ dy, const float* __restrict__ g,
                                                           const float* __restrict__ mean, const float* __restrict__ rstd, T* __restrict__ dx,
                                                           float* __restrict__ dgamma, float* __restrict__ dbeta, int C, int rows,
                                                           const T* __restrict__ dres = nullptr, float* __restrict__ dxsum = nullptr) {
  // dxsum (fp32 [C], zeroed by the caller): column sums of the dx written here = the bias gradient of the Linear whose output was
  // the `res` operand of the forward (its dy IS this dx), saving that layer a pass over dy
  // dres: gradient arriving at the normalised tensor from its other consumer (the residual path); added into dx here
  constexpr int VN = Vec16<T>::N;
  extern __shared__ float acc[];                     // [2][C], zeroed, then += the 8 warps' partial sums
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nvec = C / VN;
  const float invC = 1.f / (float)C;
  for (int c = threadIdx.x; c < 3 * C; c += blockDim.x) acc[c] = 0.f;
  float gam[kVPL][VN], dg[kVPL][VN], db[kVPL][VN], dsx[kVPL][VN];
#pragma unroll
  for (int k = 0; k < kVPL; ++k)
#pragma unroll
    for (int i = 0; i < VN; ++i) {
      const int c = (lane + 32 * k) * VN + i;
      gam[k][i] = c < C ? g[c] : 0.f; dg[k][i] = 0.f; db[k][i] = 0.f; dsx[k][i] = 0.f;
    }
  for (int r = blockIdx.x * (kT / 32) + warp; r < rows; r += gridDim.x * (kT / 32)) {
    const long o = (long)r * C;
    const float mu = mean[r], rs = rstd[r];
    float xh[kVPL][VN], d[kVPL][VN];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int k = 0; k < kVPL; ++k) {
      const int vi = lane + 32 * k;
      if (vi < nvec) {
        Vec16<T>::load(x + o + vi * VN, xh[k]);
        Vec16<T>::load(dy + o + vi * VN, d[k]);
#pragma unroll
        for (int i = 0; i < VN; ++i) {
          xh[k][i] = (xh[k][i] - mu) * rs;
          const float gy = d[k][i] * gam[k][i];
          s1 += gy; s2 += gy * xh[k][i];
          dg[k][i] += d[k][i] * xh[k][i]; db[k][i] += d[k][i];
        }
      }
    }
    s1 = warp_sum(s1) * invC; s2 = warp_sum(s2) * invC;
#pragma unroll
    for (int k = 0; k < kVPL; ++k) {
      const int vi = lane + 32 * k;
      if (vi < nvec) {
        float out[VN];
#pragma unroll
        for (int i = 0; i < VN; ++i) out[i] = rs * (d[k][i] * gam[k][i] - s1 - xh[k][i] * s2);
        if (dres != nullptr) {
          float rr[VN];
          Vec16<T>::load(dres + o + vi * VN, rr);
#pragma unroll
          for (int i = 0; i < VN; ++i) out[i] += rr[i];
        }
        if (dxsum != nullptr) {
#pragma unroll
          for (int i = 0; i < VN; ++i) dsx[k][i] += out[i];
        }
        Vec16<T>::store(dx + o + vi * VN, out);
      }
    }
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < kVPL; ++k)
    if (lane + 32 * k < nvec) {
#pragma unroll
      for (int i = 0; i < VN; ++i) {
        const int c = (lane + 32 * k) * VN + i;
        atomicAdd(&acc[c], dg[k][i]); atomicAdd(&acc[C + c], db[k][i]);
        if (dxsum != nullptr) atomicAdd(&acc[2 * C + c], dsx[k][i]);
      }
    }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    atomicAdd(&dgamma[c], acc[c]); atomicAdd(&dbeta[c], acc[C + c]);
    if (dxsum != nullptr) atomicAdd(&dxsum[c], acc[2 * C + c]);
  }
}

// four consecutive elements of a row as floats
__device__ __forceinline__ float4 ld4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ float4 ld4(const __nv_bfloat16* p) {
  const uint2 v = *reinterpret_cast<const uint2*>(p);
  const __nv_bfloat162 a = *reinterpret_cast<const __nv_bfloat162*>(&v.x), b = *reinterpret_cast<const __nv_bfloat162*>(&v.y);
  return make_float4(__low2float(a), __high2float(a), __low2float(b), __high2float(b));
}
__device__ __forceinline__ void st4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
__device__ __forceinline__ void st4(__nv_bfloat16* p, float4 v) {
  const __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y), b = __floats2bfloat162_rn(v.z, v.w);
  *reinterpret_cast<uint2*>(p) = make_uint2(*reinterpret_cast<const uint32_t*>(&a), *reinterpret_cast<const uint32_t*>(&b));
}
template <typename T> __global__ void __launch_bounds__(256) k_embed_fwd(const long* __restrict__ idx, const float* __restrict__ wte, const float* __restrict__ wpe,
                                                                         T* __restrict__ out, long n, int Tlen, int C) {
  const int c4 = C / 4;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
    const long r = i / c4; const int c = (int)(i % c4) * 4;
    const float4 a = ld4(wte + idx[r] * C + c), b = ld4(wpe + (r % Tlen) * C + c);
    st4(out + r * C + c, make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w));
  }
}
template <typename T> __global__ void __launch_bounds__(256) k_embed_bwd_wte(const long* __restrict__ idx, const T* __restrict__ dy, float* __restrict__ dwte, long n, int C) {
  const int c4 = C / 4;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
    const long r = i / c4; const int c = (int)(i % c4) * 4;
    const float4 v = ld4(dy + r * C + c);
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dwte + idx[r] * C + c), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
  }
}
template <typename T> __global__ void __launch_bounds__(128) k_embed_bwd_wpe(const T* __restrict__ dy, float* __restrict__ dwpe, long npos, int B, int Tlen, int C) {
  const int c4 = C / 4;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < npos; i += (long)gridDim.x * blockDim.x) {
    const long t = i / c4; const int c = (int)(i % c4) * 4;
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int b = 0; b < B; ++b) { const float4 v = ld4(dy + ((long)b * Tlen + t) * C + c); a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w; }
    st4(dwpe + t * C + c, a);
  }
}

__device__ __forceinline__ float gelu_tanh(float x) {
  const float k = 0.7978845608028654f, a = 0.044715f;
  return 0.5f * x * (1.f + tanhf(k * (x + a * x * x * x)));
}
__device__ __forceinline__ float gelu_tanh_grad(float x) {
  const float k = 0.7978845608028654f, a = 0.044715f;
  const float u = k * (x + a * x * x * x), t = tanhf(u);
  return 0.5f * (1.f + t) + 0.5f * x * (1.f - t * t) * k * (1.f + 3.f * a * x * x);
}
template <typename T> __global__ void k_gelu_fwd(const T* __restrict__ x, T* __restrict__ y, long n) {
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) y[i] = from_f<T>(gelu_tanh(to_f(x[i])));
}
template <typename T> __global__ void k_gelu_bwd(const T* __restrict__ x, const T* __restrict__ dy, T* __restrict__ dx, long n) {
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x)
    dx[i] = from_f<T>(to_f(dy[i]) * gelu_tanh_grad(to_f(x[i])));
}
// dx = dy * gelu'(x) AND colsum[c] += sum_r dx[r, c] (the bias gradient of the Linear that produced x): 16-byte accesses, a
// thread owns one vector of columns over a strip of rows, the eight row lanes of a CTA fold through shared memory.
template <typename T> __global__ void __launch_bounds__(256) k_gelu_bwd_colsum(const T* __restrict__ x, const T* __restrict__ dy, T* __restrict__ dx,
                                                                               float* __restrict__ colsum, long rows, int C, long rows_per_cta) {
  constexpr int VN = Vec16<T>::N;
  __shared__ float sacc[8][32 * VN + 1];
  const int lx = threadIdx.x & 31, ly = threadIdx.x >> 5;
  const long vcol = (long)blockIdx.x * 32 + lx;
  const bool live = vcol * VN < C;
  float a[VN];
#pragma unroll
  for (int i = 0; i < VN; ++i) a[i] = 0.f;
  const long r0 = (long)blockIdx.y * rows_per_cta, r1 = min(rows, r0 + rows_per_cta);
  if (live) {
    for (long r = r0 + ly; r < r1; r += 16) {
      float xv[2][VN], dv[2][VN];
      const bool two = r + 8 < r1;
      Vec16<T>::load(x + r * C + vcol * VN, xv[0]); Vec16<T>::load(dy + r * C + vcol * VN, dv[0]);
      if (two) { Vec16<T>::load(x + (r + 8) * C + vcol * VN, xv[1]); Vec16<T>::load(dy + (r + 8) * C + vcol * VN, dv[1]); }
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        if (u == 1 && !two) break;
        float o[VN];
#pragma unroll
        for (int i = 0; i < VN; ++i) { o[i] = to_f(from_f<T>(dv[u][i] * gelu_tanh_grad(xv[u][i]))); a[i] += o[i]; }
        Vec16<T>::store(dx + (r + 8 * u) * C + vcol * VN, o);
      }
    }
  }
#pragma unroll
  for (int i = 0; i < VN; ++i) sacc[ly][lx * VN + i] = a[i];
  __syncthreads();
  for (int i = threadIdx.x; i < 32 * VN; i += 256) {
    const long c = (long)blockIdx.x * 32 * VN + i;
    if (c < C) {
      float t = 0.f;
#pragma unroll
      for (int k = 0; k < 8; ++k) t += sacc[k][i];
      atomicAdd(&colsum[c], t);
    }
  }
}
}  // namespace

extern "C" {
int dcp_layernorm_fwd(const void* x, void* y, const float* gamma, const float* beta, float* mean, float* rstd, int rows, int C, float eps, int dtype,
                      cudaStream_t s) {
  const int vn = dtype == DCP_F32 ? 4 : 8;
  if (C % vn == 0 && C / vn <= 32 * kMaxVPL && ((uintptr_t)x % 16 == 0) && ((uintptr_t)y % 16 == 0)) {
    const int ctas = min((rows + kT / 32 - 1) / (kT / 32), 148 * 8);
    const int vpl = (C / vn + 31) / 32;
#define LN_FWD(V) DISPATCH_ACT(dtype, (k_layernorm_fwd_warp<T, V><<<ctas, kT, 0, s>>>((const T*)x, (T*)y, gamma, beta, mean, rstd, C, rows, eps)))
    if (vpl == 1) { LN_FWD(1); } else if (vpl == 2) { LN_FWD(2); } else if (vpl == 3) { LN_FWD(3); } else { LN_FWD(4); }
#undef LN_FWD
    DCP_CUDA_CHECK(cudaGetLastError()); return 0;
  }
  DISPATCH_ACT(dtype, (k_layernorm_fwd<T><<<rows, kT, 0, s>>>((const T*)x, (T*)y, gamma, beta, mean, rstd, C, eps)));
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}
// dgamma / dbeta: fp32 [C], zeroed here.
int dcp_layernorm_bwd(const void* x, const void* dy, const float* gamma, const float* mean, const float* rstd, void* dx, float* dgamma, float* dbeta,
                      int rows, int C, int dtype, cudaStream_t s) {
  DCP_CUDA_CHECK(cudaMemsetAsync(dgamma, 0, sizeof(float) * C, s));
  DCP_CUDA_CHECK(cudaMemsetAsync(dbeta, 0, sizeof(float) * C, s));
  const int vn = dtype == DCP_F32 ? 4 : 8;
  if (C % vn == 0 && C / vn <= 32 * kMaxVPL && ((uintptr_t)x % 16 == 0) && ((uintptr_t)dy % 16 == 0) && ((uintptr_t)dx % 16 == 0)) {
    const int ctasw = min((rows + kT / 32 - 1) / (kT / 32), 148 * 2);
    const int vpl = (C / vn + 31) / 32;
#define LN_BWD(V) DISPATCH_ACT(dtype, (k_layernorm_bwd_warp<T, V><<<ctasw, kT, sizeof(float) * 3 * C, s>>>((const T*)x, (const T*)dy, gamma, mean, rstd, (T*)dx, dgamma, dbeta, C, rows)))
    if (vpl == 1) { LN_BWD(1); } else if (vpl == 2) { LN_BWD(2); } else if (vpl == 3) { LN_BWD(3); } else { LN_BWD(4); }
#undef LN_BWD
    DCP_CUDA_CHECK(cudaGetLastError()); return 0;
  }
  int ctas = rows < 148 * 2 ? rows : 148 * 2;
  int rpc = (rows + ctas - 1) / ctas;
  ctas = (rows + rpc - 1) / rpc;
  DISPATCH_ACT(dtype, (k_layernorm_bwd<T><<<ctas, kT, sizeof(float) * 2 * C, s>>>((const T*)x, (const T*)dy, gamma, mean, rstd, (T*)dx, dgamma, dbeta, C, rows, rpc)));
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}
// y = LayerNorm(x + res), sum_out = x + res (the residual stream after the add): one pass instead of an add kernel plus a
// LayerNorm.  Row length as in the warp-per-row path of dcp_layernorm_fwd (C a multiple of the 16-byte vector, <= 1024 bf16 /
// 512 fp32 elements), 16-byte aligned pointers; returns -1 otherwise (the caller falls back to add + layer norm).
int dcp_add_layernorm_fwd(const void* x, const void* res, void* sum_out, void* y, const float* gamma, const float* beta, float* mean, float* rstd,
                          int rows, int C, float eps, int dtype, cudaStream_t s) {
  const int vn = dtype == DCP_F32 ? 4 : 8;
  if (!(C % vn == 0 && C / vn <= 32 * kMaxVPL) || ((uintptr_t)x | (uintptr_t)res | (uintptr_t)sum_out | (uintptr_t)y) % 16 != 0) {
    dcp_set_error("add_layernorm: unsupported row length %d or unaligned pointers", C); return -1;
  }
  const int ctas = min((rows + kT / 32 - 1) / (kT / 32), 148 * 8);
  const int vpl = (C / vn + 31) / 32;
#define LN_FWD(V) DISPATCH_ACT(dtype, (k_layernorm_fwd_warp<T, V><<<ctas, kT, 0, s>>>((const T*)x, (T*)y, gamma, beta, mean, rstd, C, rows, eps, (const T*)res, (T*)sum_out)))
  if (vpl == 1) { LN_FWD(1); } else if (vpl == 2) { LN_FWD(2); } else if (vpl == 3) { LN_FWD(3); } else { LN_FWD(4); }
#undef LN_FWD
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}
// Backward of dcp_add_layernorm_fwd: dx = LayerNorm-backward(dy) + dsum, where dsum is the gradient that reached the sum from
// the residual path (the same dx is the gradient of both x and res).  x = the saved sum.  dgamma / dbeta: fp32 [C], zeroed here.
// dxsum: optional fp32 [C] that receives the column sums of dx (zeroed here) - the bias gradient of the Linear that produced res.
int dcp_add_layernorm_bwd(const void* x, const void* dy, const void* dsum, const float* gamma, const float* mean, const float* rstd, void* dx,
                          float* dgamma, float* dbeta, float* dxsum, int rows, int C, int dtype, cudaStream_t s) {
  const int vn = dtype == DCP_F32 ? 4 : 8;
  if (!(C % vn == 0 && C / vn <= 32 * kMaxVPL) || ((uintptr_t)x | (uintptr_t)dy | (uintptr_t)dsum | (uintptr_t)dx) % 16 != 0) {
    dcp_set_error("add_layernorm backward: unsupported row length %d or unaligned pointers", C); return -1;
  }
  DCP_CUDA_CHECK(cudaMemsetAsync(dgamma, 0, sizeof(float) * C, s));
  DCP_CUDA_CHECK(cudaMemsetAsync(dbeta, 0, sizeof(float) * C, s));
  if (dxsum != nullptr) DCP_CUDA_CHECK(cudaMemsetAsync(dxsum, 0, sizeof(float) * C, s));
  const int ctasw = min((rows + kT / 32 - 1) / (kT / 32), 148 * 2);
  const int vpl = (C / vn + 31) / 32;
#define LN_BWD(V) DISPATCH_ACT(dtype, (k_layernorm_bwd_warp<T, V><<<ctasw, kT, sizeof(float) * 3 * C, s>>>((const T*)x, (const T*)dy, gamma, mean, rstd, (T*)dx, dgamma, dbeta, C, rows, (const T*)dsum, dxsum)))
  if (vpl == 1) { LN_BWD(1); } else if (vpl == 2) { LN_BWD(2); } else if (vpl == 3) { LN_BWD(3); } else { LN_BWD(4); }
#undef LN_BWD
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}

// Token + position embedding of a transformer: out[r, :] = T(wte[idx[r], :] + wpe[r % Tlen, :]) for r < rows = B * Tlen.
// fp32 tables, f32 / bf16 output, C a multiple of 4.  (One gather kernel instead of two gathers, an add and a cast.)
int dcp_embed_fwd(const long* idx, const float* wte, const float* wpe, void* out, long rows, int Tlen, int C, int dtype, cudaStream_t s) {
  if (C % 4) { dcp_set_error("embed: C=%d must be a multiple of 4", C); return -1; }
  const long n = rows * (C / 4);
  DISPATCH_ACT(dtype, (k_embed_fwd<T><<<grid_1d(n, 256), 256, 0, s>>>(idx, wte, wpe, (T*)out, n, Tlen, C)));
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}
// Backward: dwte[idx[r], :] += dy[r, :] (vector reductions; dwte fp32 [V, C] zeroed by the caller - rows that are no token stay
// zero) and dwpe[t, :] = sum_b dy[b * Tlen + t, :] for t < Tlen (plain stores; the caller zeroes the rows beyond Tlen).
int dcp_embed_bwd(const long* idx, const void* dy, float* dwte, float* dwpe, long rows, int Tlen, int C, int dtype, cudaStream_t s) {
  if (C % 4) { dcp_set_error("embed: C=%d must be a multiple of 4", C); return -1; }
  const long n = rows * (C / 4), npos = (long)Tlen * (C / 4);
  DISPATCH_ACT(dtype, (k_embed_bwd_wte<T><<<grid_1d(n, 256), 256, 0, s>>>(idx, (const T*)dy, dwte, n, C)));
  DISPATCH_ACT(dtype, (k_embed_bwd_wpe<T><<<grid_1d(npos, 128), 128, 0, s>>>((const T*)dy, dwpe, npos, (int)(rows / Tlen), Tlen, C)));
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}
// GELU backward over a [rows, C] matrix that also produces the column sums of dx (fp32 [C], zeroed here): C a multiple of the
// 16-byte vector, aligned pointers (-1 otherwise; the caller falls back to dcp_gelu_bwd).
int dcp_gelu_bwd_colsum(const void* x, const void* dy, void* dx, float* colsum, long rows, int C, int dtype, cudaStream_t s) {
  const int vn = dtype == DCP_F32 ? 4 : 8;
  if (C % vn || ((uintptr_t)x | (uintptr_t)dy | (uintptr_t)dx) % 16 != 0) { dcp_set_error("gelu_bwd_colsum: C=%d / alignment unsupported", C); return -1; }
  DCP_CUDA_CHECK(cudaMemsetAsync(colsum, 0, sizeof(float) * C, s));
  const int gx = (C / vn + 31) / 32;
  long gy = (148L * 4 + gx - 1) / gx; const long maxy = (rows + 15) / 16;
  if (gy > maxy) gy = maxy; if (gy < 1) gy = 1;
  const long rpc = ((rows + gy - 1) / gy + 15) / 16 * 16;          // strips start on multiples of 16 rows
  gy = (rows + rpc - 1) / rpc;
  DISPATCH_ACT(dtype, (k_gelu_bwd_colsum<T><<<dim3(gx, (unsigned)gy), 256, 0, s>>>((const T*)x, (const T*)dy, (T*)dx, colsum, rows, C, rpc)));
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}
int dcp_gelu_fwd(const void* x, void* y, long n, int dtype, cudaStream_t s) {
  DISPATCH_ACT(dtype, (k_gelu_fwd<T><<<grid_1d(n, 256), 256, 0, s>>>((const T*)x, (T*)y, n)));
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}
int dcp_gelu_bwd(const void* x, const void* dy, void* dx, long n, int dtype, cudaStream_t s) {
  DISPATCH_ACT(dtype, (k_gelu_bwd<T><<<grid_1d(n, 256), 256, 0, s>>>((const T*)x, (const T*)dy, (T*)dx, n)));
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}
}  // extern "C"
