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

// dz1 [P,32], x [N,28,28] -> dw1[32][9] += , db1[32] +=   (8 pixel lanes x 32 channels per CTA; each lane keeps
// 4 pixels in flight so the dependent-load chain is short; one atomic per output per CTA)
__global__ void __launch_bounds__(256) k_conv1_wgrad(const float* __restrict__ dz1, const float* __restrict__ x, float* __restrict__ dw1,
                                                     float* __restrict__ db1, long total_pixels) {
  __shared__ float red[8][32][11];
  const int c = threadIdx.x & 31, lane = threadIdx.x >> 5;
  float acc[10];
#pragma unroll
  for (int t = 0; t < 10; ++t) acc[t] = 0.f;
  const long stride = (long)gridDim.x * 8;
  for (long p0 = (long)blockIdx.x * 8 + lane; p0 < total_pixels; p0 += 4 * stride) {
    float g[4]; const float* xi[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long pix = p0 + u * stride;
      const bool ok = pix < total_pixels;
      g[u] = ok ? dz1[pix * 32 + c] : 0.f;
      const long pc = ok ? pix : 0;
      const int n = (int)(pc / P1), q = (int)(pc % P1), h = q / HW1, wq = q % HW1;
      xi[u] = x + (long)n * 784 + h * 28 + wq;
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
#pragma unroll
      for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int s = 0; s < 3; ++s) acc[r * 3 + s] = fmaf(g[u], __ldg(xi[u] + r * 28 + s), acc[r * 3 + s]);
      acc[9] += g[u];
    }
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
  k_conv1_wgrad<<<148 * 2, 256, 0, s>>>(dz1, x, dw1, db1, (long)N * P1);
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}

}  // extern "C"

// =====================================================================================================
// Head of the network in ONE kernel, forward AND backward (K9-K14 and their gradients, ref:main.py:40-44,61):
//   h1[N,128] -> BatchNorm1d (batch statistics, running-stat update) -> ReLU -> dropout p2 -> fc2 (128 -> 10)
//   -> log_softmax + nll (mean) -> d logits -> fc2 weight/bias/data gradients -> dropout / ReLU backward
//   -> BatchNorm backward -> dh1[N,128], dgamma, dbeta.
// The whole problem is 16 K activations, so it lives in the shared memory of a single 1024-thread CTA; the
// 13 launches it replaces cost ~65 us of launch latency and tail effects.  N <= 128 rows.
// Thread (f = tid % 128, rg = tid / 128) owns feature f of rows rg, rg + 8, ...
// =====================================================================================================
namespace {
constexpr int HF = 128, HC = 10, HRG = 8, HMAXN = 128, HR = HMAXN / HRG;

struct HeadSmem {
  float xhat[HMAXN][HF];
  float h3[HMAXN][HF];
  float w2[HC][HF];
  float dlog[HMAXN][HC + 2];
  float red[HRG][HF][HC + 1];
  float stat[4][HF];      // mean, rstd, dbeta/N, dgamma/N
  float loss;
};

__global__ void __launch_bounds__(1024, 1)
k_convnet_head(const float* __restrict__ h1, const long long* __restrict__ target, const float* __restrict__ gamma, const float* __restrict__ beta,
               float* __restrict__ run_mean, float* __restrict__ run_var, long long* __restrict__ nbt, const float* __restrict__ w2,
               const float* __restrict__ b2, float* __restrict__ loss_out, float* __restrict__ logits_out, float* __restrict__ dh1,
               float* __restrict__ dgamma, float* __restrict__ dbeta, float* __restrict__ dw2, float* __restrict__ db2, int N, float eps,
               float momentum, float p2, unsigned long long seed, unsigned long long stream, const long long* __restrict__ step, int training) {
  extern __shared__ __align__(16) uint8_t head_raw[];
  HeadSmem& S = *reinterpret_cast<HeadSmem*>(head_raw);
  const int tid = threadIdx.x, f = tid & (HF - 1), rg = tid >> 7, warp = tid >> 5, lane = tid & 31;
  const float invN = 1.f / (float)N;
  float x[HR];
  // ---- 1. batch statistics (two-pass, data stays in registers) ----
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < HR; ++i) { const int r = rg + i * HRG; x[i] = r < N ? h1[r * HF + f] : 0.f; s += x[i]; }
  S.red[rg][f][0] = s;
  for (int i = tid; i < HC * HF; i += 1024) S.w2[i / HF][i % HF] = w2[i];
  if (tid == 0) S.loss = 0.f;
  __syncthreads();
  if (rg == 0) { float t = 0.f; for (int g = 0; g < HRG; ++g) t += S.red[g][f][0]; S.stat[0][f] = t * invN; }
  __syncthreads();
  const float mean = S.stat[0][f];
  float v = 0.f;
#pragma unroll
  for (int i = 0; i < HR; ++i) { const int r = rg + i * HRG; if (r < N) { const float d = x[i] - mean; v += d * d; } }
  S.red[rg][f][1] = v;
  __syncthreads();
  if (rg == 0) {
    float t = 0.f; for (int g = 0; g < HRG; ++g) t += S.red[g][f][1];
    const float var = t * invN;
    S.stat[1][f] = rsqrtf(var + eps);
    if (training && run_mean) {
      const float unbiased = N > 1 ? var * (float)N / (float)(N - 1) : var;
      run_mean[f] = (1.f - momentum) * run_mean[f] + momentum * mean;
      run_var[f] = (1.f - momentum) * run_var[f] + momentum * unbiased;
    }
  }
  if (tid == 0 && training && nbt) *nbt += 1;
  __syncthreads();
  // ---- 2. normalise + affine + ReLU + dropout ----
  const float rstd = S.stat[1][f], gm = gamma[f], bt = beta[f];
  const unsigned long long st = stream + (step ? (unsigned long long)(*step) * 0x10000ull : 0ull);
  const float dscale = (training && p2 > 0.f) ? 1.f / (1.f - p2) : 1.f;
  unsigned keep = 0, pos = 0;   // bit i: dropout kept / ReLU active for row rg + 8 i
#pragma unroll
  for (int i = 0; i < HR; ++i) {
    const int r = rg + i * HRG;
    float xh = 0.f, h = 0.f;
    if (r < N) {
      xh = (x[i] - mean) * rstd;
      float a = xh * gm + bt;
      if (a > 0.f) pos |= 1u << i; else a = 0.f;
      bool k = true;
      if (training && p2 > 0.f) k = philox_uniform(seed, st, (unsigned long long)(r * HF + f)) >= p2;
      if (k) keep |= 1u << i;
      h = k ? a * dscale : 0.f;
    }
    S.xhat[r][f] = xh;      // r < HMAXN always: rg < 8, i < 16
    S.h3[r][f] = h;
  }
  __syncthreads();
  // ---- 3. fc2 + log-softmax + nll + dlogits: one warp per row ----
  for (int r = warp; r < N; r += 32) {
    float acc[HC];
#pragma unroll
    for (int j = 0; j < HC; ++j) acc[j] = 0.f;
#pragma unroll
    for (int q = 0; q < HF / 32; ++q) {
      const float hv = S.h3[r][lane + 32 * q];
#pragma unroll
      for (int j = 0; j < HC; ++j) acc[j] = fmaf(hv, S.w2[j][lane + 32 * q], acc[j]);
    }
#pragma unroll
    for (int j = 0; j < HC; ++j) acc[j] = warp_sum(acc[j]) + b2[j];
    float m = acc[0];
#pragma unroll
    for (int j = 1; j < HC; ++j) m = fmaxf(m, acc[j]);
    float se = 0.f;
#pragma unroll
    for (int j = 0; j < HC; ++j) se += __expf(acc[j] - m);
    const float lse = m + __logf(se);
    const int t = (int)target[r];
    if (lane < HC) {
      const float pj = __expf(acc[lane] - lse);
      S.dlog[r][lane] = (pj - (lane == t ? 1.f : 0.f)) * invN;
      if (logits_out) logits_out[r * HC + lane] = acc[lane] - lse;     // log-probabilities, as the model's forward returns
    }
    if (lane == 0 && t >= 0 && t < HC) atomicAdd(&S.loss, (lse - acc[t]) * invN);
  }
  __syncthreads();
  if (tid == 0) *loss_out = S.loss;
  // ---- 4. fc2 gradients ----
  float gw[HC];
#pragma unroll
  for (int j = 0; j < HC; ++j) gw[j] = 0.f;
#pragma unroll
  for (int i = 0; i < HR; ++i) {
    const int r = rg + i * HRG;
    if (r < N) {
      const float hv = S.h3[r][f];
#pragma unroll
      for (int j = 0; j < HC; ++j) gw[j] = fmaf(S.dlog[r][j], hv, gw[j]);
    }
  }
#pragma unroll
  for (int j = 0; j < HC; ++j) S.red[rg][f][j] = gw[j];
  __syncthreads();
  for (int i = tid; i < HC * HF; i += 1024) {
    const int j = i / HF, ff = i % HF;
    float t = 0.f;
#pragma unroll
    for (int g = 0; g < HRG; ++g) t += S.red[g][ff][j];
    dw2[i] = t;
  }
  if (tid < HC) { float t = 0.f; for (int r = 0; r < N; ++r) t += S.dlog[r][tid]; db2[tid] = t; }
  // ---- 5. dh3 -> dropout / ReLU backward -> BatchNorm backward ----
  float g[HR];
  float sb = 0.f, sg = 0.f;
#pragma unroll
  for (int i = 0; i < HR; ++i) {
    const int r = rg + i * HRG;
    float gv = 0.f;
    if (r < N && ((keep >> i) & 1u) && ((pos >> i) & 1u)) {
#pragma unroll
      for (int j = 0; j < HC; ++j) gv = fmaf(S.dlog[r][j], S.w2[j][f], gv);
      gv *= dscale;
    }
    g[i] = gv;
    sb += gv;
    sg += gv * S.xhat[r][f];
  }
  __syncthreads();                       // S.red is reused
  S.red[rg][f][0] = sb; S.red[rg][f][1] = sg;
  __syncthreads();
  if (rg == 0) {
    float tb = 0.f, tg = 0.f;
    for (int q = 0; q < HRG; ++q) { tb += S.red[q][f][0]; tg += S.red[q][f][1]; }
    dbeta[f] = tb; dgamma[f] = tg;
    S.stat[2][f] = tb * invN; S.stat[3][f] = tg * invN;
  }
  __syncthreads();
  const float k1 = gm * rstd, k2 = S.stat[2][f], k3 = S.stat[3][f];
#pragma unroll
  for (int i = 0; i < HR; ++i) {
    const int r = rg + i * HRG;
    if (r < N) dh1[r * HF + f] = k1 * (g[i] - k2 - S.xhat[r][f] * k3);
  }
}
}  // namespace

extern "C" {
// Fused head (see above).  N <= 128.  logits_out (log-probabilities [N,10]) may be null.  All gradient outputs are
// fully overwritten.  Training mode only (batch statistics); p2 = dropout probability of the second dropout.
int dcp_convnet_head(const float* h1, const void* target, const float* gamma, const float* beta, float* run_mean, float* run_var, void* nbt,
                     const float* w2, const float* b2, float* loss, float* logits_out, float* dh1, float* dgamma, float* dbeta, float* dw2,
                     float* db2, int N, float eps, float momentum, float p2, unsigned long long seed, unsigned long long stream_id,
                     const void* step, int training, cudaStream_t s) {
  if (N < 1 || N > HMAXN) { dcp_set_error("convnet head: batch %d not in [1, %d]", N, HMAXN); return -1; }
  static bool attr = false;
  if (!attr) { DCP_CUDA_CHECK(cudaFuncSetAttribute(k_convnet_head, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(HeadSmem))); attr = true; }
  k_convnet_head<<<1, 1024, sizeof(HeadSmem), s>>>(h1, (const long long*)target, gamma, beta, run_mean, run_var, (long long*)nbt, w2, b2, loss,
                                                   logits_out, dh1, dgamma, dbeta, dw2, db2, N, eps, momentum, p2, seed, stream_id,
                                                   (const long long*)step, training);
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}
}  // extern "C"
