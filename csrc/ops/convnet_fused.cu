// Fused kernels of the flagship training step (the reference's ConvNet, ref:main.py:20-45) that work on the
// "flat pixel" geometry used by conv_flat_tc.cu: every activation of the convolutional trunk is a matrix
// [N * 26 * 26, C] in the geometry of conv1's output; conv2's 24x24 output and its gradient live in the same
// 26x26 frame with zeros outside the valid window.
//
//   conv1_fwd      x[N,28,28] -> a1[N*676, 32] = relu(conv3x3 + b)                   (K1+K2, ref:main.py:32-33)
//   pool_fwd       a2[N*676, 64] -> pooled[N, 64*12*12] (NCHW flatten order) with 2x2 max-pool, channel
//                  dropout p (Philox, regenerated in backward) and the argmax byte     (K5-K7, ref:main.py:36-38)
//   pool_bwd       dpooled -> dy2[N*676, 64]: routes the gradient to the argmax position, applies the dropout
//                  mask and the ReLU mask (pooled > 0), zero elsewhere                 (backward of K4-K6)
//   conv1_wgrad    dz1[N*676, 32], x -> dw1[32, 9], db1[32] (fp32 atomics; caller zeroes)   (backward of K1)
#include "ops/ops_common.cuh"

namespace {

constexpr int HW1 = 26, P1 = 676;   // conv1 output frame
constexpr int HP = 12;              // pooled extent

__global__ void __launch_bounds__(256) k_conv1_fwd(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ b,
                                                   float* __restrict__ a1, long total_pixels) {
  __shared__ float sw[32 * 9], sb[32];
  for (int i = threadIdx.x; i < 32 * 9; i += blockDim.x) sw[i] = w[i];
  if (threadIdx.x < 32) sb[threadIdx.x] = b[threadIdx.x];
  __syncthreads();
  // 4 threads per pixel, 8 channels each
  const long gid = (long)blockIdx.x * blockDim.x + threadIdx.x;
  const long pix = gid >> 2;
  const int cg = (int)(gid & 3) * 8;
  if (pix >= total_pixels) return;
  const int n = (int)(pix / P1), q = (int)(pix % P1), h = q / HW1, wq = q % HW1;
  const float* xi = x + (long)n * 784 + h * 28 + wq;
  float v[9];
#pragma unroll
  for (int r = 0; r < 3; ++r)
#pragma unroll
    for (int s = 0; s < 3; ++s) v[r * 3 + s] = __ldg(xi + r * 28 + s);
  float o[8];
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    float acc = sb[cg + c];
#pragma unroll
    for (int t = 0; t < 9; ++t) acc = fmaf(v[t], sw[(cg + c) * 9 + t], acc);
    o[c] = fmaxf(acc, 0.f);
  }
  float4* dst = reinterpret_cast<float4*>(a1 + pix * 32 + cg);
  dst[0] = make_float4(o[0], o[1], o[2], o[3]);
  dst[1] = make_float4(o[4], o[5], o[6], o[7]);
}

// grid (12 pooled rows, N); 256 threads.  a2 rows 2*ph, 2*ph+1 of image n -> pooled[n][c][ph][0..11]
__global__ void __launch_bounds__(256) k_pool_fwd(const float* __restrict__ a2, float* __restrict__ pooled, uint8_t* __restrict__ arg, float p,
                                                  unsigned long long seed, unsigned long long stream, const long long* __restrict__ step, int training) {
  __shared__ float sv[HP][65];
  __shared__ uint8_t sa[HP][65];
  const int ph = blockIdx.x, n = blockIdx.y;
  const float* base = a2 + ((long)n * P1 + (long)(2 * ph) * HW1) * 64;
  for (int i = threadIdx.x; i < HP * 64; i += blockDim.x) {
    const int c = i & 63, pw = i >> 6;
    const float* q = base + (2 * pw) * 64 + c;
    float v0 = q[0], v1 = q[64], v2 = q[HW1 * 64], v3 = q[HW1 * 64 + 64];
    float m = v0; int a = 0;
    if (v1 > m) { m = v1; a = 1; }
    if (v2 > m) { m = v2; a = 2; }
    if (v3 > m) { m = v3; a = 3; }
    sv[pw][c] = m; sa[pw][c] = (uint8_t)a;
  }
  __syncthreads();
  const unsigned long long st = stream + (step ? (unsigned long long)(*step) * 0x10000ull : 0ull);
  const float scale = 1.f / (1.f - p);
  for (int i = threadIdx.x; i < 64 * HP; i += blockDim.x) {
    const int pw = i % HP, c = i / HP;
    float m = sv[pw][c];
    if (training && p > 0.f) {
      float u = philox_uniform(seed, st, (unsigned long long)(n * 64 + c));
      m = u >= p ? m * scale : 0.f;
    }
    const long o = (((long)n * 64 + c) * HP + ph) * HP + pw;
    pooled[o] = m;
    arg[o] = sa[pw][c];
  }
}

// grid (13, N): block ph < 12 writes rows 2ph, 2ph+1 of dy2 (26 pixels wide, zeros at w >= 24); block 12 zeroes rows 24, 25.
__global__ void __launch_bounds__(256) k_pool_bwd(const float* __restrict__ dpooled, const float* __restrict__ pooled, const uint8_t* __restrict__ arg,
                                                  float* __restrict__ dy2, float p, unsigned long long seed, unsigned long long stream,
                                                  const long long* __restrict__ step, int training) {
  __shared__ float sg[HP][65];
  __shared__ uint8_t sa[HP][65];
  const int ph = blockIdx.x, n = blockIdx.y;
  float* out = dy2 + ((long)n * P1 + (long)(2 * ph) * HW1) * 64;
  if (ph == HP) {
    float4* o4 = reinterpret_cast<float4*>(out);
    for (int i = threadIdx.x; i < 2 * HW1 * 64 / 4; i += blockDim.x) o4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    return;
  }
  const unsigned long long st = stream + (step ? (unsigned long long)(*step) * 0x10000ull : 0ull);
  const float scale = 1.f / (1.f - p);
  for (int i = threadIdx.x; i < 64 * HP; i += blockDim.x) {
    const int pw = i % HP, c = i / HP;
    const long o = (((long)n * 64 + c) * HP + ph) * HP + pw;
    float g = dpooled[o];
    const float y = pooled[o];
    if (!(y > 0.f)) g = 0.f;                       // ReLU mask (a dropped channel has pooled == 0 as well)
    else if (training && p > 0.f) g *= scale;      // kept channel: the forward multiplied by 1/(1-p)
    sg[pw][c] = g; sa[pw][c] = arg[o];
  }
  __syncthreads();
  (void)seed; (void)st;
  for (int i = threadIdx.x; i < 2 * HW1 * 64; i += blockDim.x) {
    const int c = i & 63, px = (i >> 6) % HW1, row = (i >> 6) / HW1;
    float v = 0.f;
    if (px < 2 * HP) {
      const int pw = px >> 1, pos = row * 2 + (px & 1);
      if (sa[pw][c] == pos) v = sg[pw][c];
    }
    out[i] = v;
  }
}

// dz1 [P,32], x [N,28,28] -> dw1[32][9] += , db1[32] +=   (8 pixel lanes x 32 channels per CTA)
__global__ void __launch_bounds__(256) k_conv1_wgrad(const float* __restrict__ dz1, const float* __restrict__ x, float* __restrict__ dw1,
                                                     float* __restrict__ db1, long total_pixels) {
  __shared__ float red[8][32][11];
  const int c = threadIdx.x & 31, lane = threadIdx.x >> 5;
  float acc[10];
#pragma unroll
  for (int t = 0; t < 10; ++t) acc[t] = 0.f;
  for (long pix = (long)blockIdx.x * 8 + lane; pix < total_pixels; pix += (long)gridDim.x * 8) {
    const float g = dz1[pix * 32 + c];
    const int n = (int)(pix / P1), q = (int)(pix % P1), h = q / HW1, wq = q % HW1;
    const float* xi = x + (long)n * 784 + h * 28 + wq;
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
      for (int s = 0; s < 3; ++s) acc[r * 3 + s] = fmaf(g, __ldg(xi + r * 28 + s), acc[r * 3 + s]);
    acc[9] += g;
  }
#pragma unroll
  for (int t = 0; t < 10; ++t) red[lane][c][t] = acc[t];
  __syncthreads();
  for (int i = threadIdx.x; i < 32 * 10; i += blockDim.x) {
    const int cc = i / 10, t = i % 10;
    float s = 0.f;
#pragma unroll
    for (int l = 0; l < 8; ++l) s += red[l][cc][t];
    if (t < 9) atomicAdd(&dw1[cc * 9 + t], s); else atomicAdd(&db1[cc], s);
  }
}

}  // namespace

extern "C" {

int dcp_convnet_conv1_fwd(const float* x, const float* w, const float* b, float* a1, int N, cudaStream_t s) {
  const long pixels = (long)N * P1;
  k_conv1_fwd<<<(unsigned)((pixels * 4 + 255) / 256), 256, 0, s>>>(x, w, b, a1, pixels);
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}
int dcp_convnet_pool_fwd(const float* a2, float* pooled, void* arg, int N, float p, unsigned long long seed, unsigned long long stream,
                         const void* step, int training, cudaStream_t s) {
  k_pool_fwd<<<dim3(HP, N), 256, 0, s>>>(a2, pooled, (uint8_t*)arg, p, seed, stream, (const long long*)step, training);
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}
int dcp_convnet_pool_bwd(const float* dpooled, const float* pooled, const void* arg, float* dy2, int N, float p, unsigned long long seed,
                         unsigned long long stream, const void* step, int training, cudaStream_t s) {
  k_pool_bwd<<<dim3(HP + 1, N), 256, 0, s>>>(dpooled, pooled, (const uint8_t*)arg, dy2, p, seed, stream, (const long long*)step, training);
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}
// dw1 / db1 are accumulated into: the caller zeroes them (they normally live in the gradient bucket).
int dcp_convnet_conv1_wgrad(const float* dz1, const float* x, float* dw1, float* db1, int N, cudaStream_t s) {
  k_conv1_wgrad<<<148 * 4, 256, 0, s>>>(dz1, x, dw1, db1, (long)N * P1);
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}

}  // extern "C"
