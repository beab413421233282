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
}  // namespace

extern "C" {
int dcp_layernorm_fwd(const void* x, void* y, const float* gamma, const float* beta, float* mean, float* rstd, int rows, int C, float eps, int dtype,
                      cudaStream_t s) {
  DISPATCH_ACT(dtype, (k_layernorm_fwd<T><<<rows, kT, 0, s>>>((const T*)x, (T*)y, gamma, beta, mean, rstd, C, eps)));
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}
// dgamma / dbeta: fp32 [C], zeroed here.
int dcp_layernorm_bwd(const void* x, const void* dy, const float* gamma, const float* mean, const float* rstd, void* dx, float* dgamma, float* dbeta,
                      int rows, int C, int dtype, cudaStream_t s) {
  DCP_CUDA_CHECK(cudaMemsetAsync(dgamma, 0, sizeof(float) * C, s));
  DCP_CUDA_CHECK(cudaMemsetAsync(dbeta, 0, sizeof(float) * C, s));
  int ctas = rows < 148 * 2 ? rows : 148 * 2;
  int rpc = (rows + ctas - 1) / ctas;
  ctas = (rows + rpc - 1) / rpc;
  DISPATCH_ACT(dtype, (k_layernorm_bwd<T><<<ctas, kT, sizeof(float) * 2 * C, s>>>((const T*)x, (const T*)dy, gamma, mean, rstd, (T*)dx, dgamma, dbeta, C, rows, rpc)));
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
