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
//   conv1_wgrad    dz1[N*676, 32], x -> dw1[32, 9], db1[32] (per-CTA partials, last CTA sums)  (backward of K1)
#include <cstdlib>

#include "ops/dp_push.cuh"
#include "ops/ops_common.cuh"
#include "ops/optim_math.cuh"

namespace {

constexpr int HW1 = 26, P1 = 676;   // conv1 output frame
constexpr int HP = 12;              // pooled extent

__global__ void __launch_bounds__(256) k_conv1_fwd(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ b,
                                                   float* __restrict__ a1, long total_pixels) {
  __shared__ float sw[32 * 9], sb[32];
  griddep_launch();
  griddep_wait();
  for (int i = threadIdx.x; i < 32 * 9; i += blockDim.x) sw[i] = w[i];
  if (threadIdx.x < 32) sb[threadIdx.x] = b[threadIdx.x];
  __syncthreads();
  // 2 threads per pixel, 16 channels each: the 9 input taps are loaded once per 16 outputs and every thread writes 64 contiguous bytes
  const long gid = (long)blockIdx.x * blockDim.x + threadIdx.x;
  const long pix = gid >> 1;
  const int cg = (int)(gid & 1) * 16;
  if (pix >= total_pixels) return;
  const int n = (int)(pix / P1), q = (int)(pix % P1), h = q / HW1, wq = q % HW1;
  const float* xi = x + (long)n * 784 + h * 28 + wq;
  float v[9];
#pragma unroll
  for (int r = 0; r < 3; ++r)
#pragma unroll
    for (int s = 0; s < 3; ++s) v[r * 3 + s] = __ldg(xi + r * 28 + s);
  float4* dst = reinterpret_cast<float4*>(a1 + pix * 32 + cg);
#pragma unroll
  for (int c4 = 0; c4 < 4; ++c4) {
    float o[4];
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      float acc = sb[cg + c4 * 4 + c];
#pragma unroll
      for (int t = 0; t < 9; ++t) acc = fmaf(v[t], sw[(cg + c4 * 4 + c) * 9 + t], acc);
      o[c] = fmaxf(acc, 0.f);
    }
    dst[c4] = make_float4(o[0], o[1], o[2], o[3]);
  }
}

// grid (12 pooled rows, N); 256 threads.  a2 rows 2*ph, 2*ph+1 of image n -> pooled[n][c][ph][0..11]
__global__ void __launch_bounds__(256) k_pool_fwd(const float* __restrict__ a2, float* __restrict__ pooled, uint8_t* __restrict__ arg, float p,
                                                  unsigned long long seed, unsigned long long stream, const long long* __restrict__ step, int training) {
  __shared__ float sv[HP][65];
  __shared__ uint8_t sa[HP][65];
  __shared__ float keep[64];          // dropout factor of channel c of this image: 0 or 1/(1-p)
  griddep_launch();
  griddep_wait();
  const int ph = blockIdx.x, n = blockIdx.y;
  const float* base = a2 + ((long)n * P1 + (long)(2 * ph) * HW1) * 64;
  if (threadIdx.x < 64) {
    float k = 1.f;
    if (training && p > 0.f) {
      const unsigned long long st = stream + (step ? (unsigned long long)(*step) * 0x10000ull : 0ull);
      k = philox_uniform(seed, st, (unsigned long long)(n * 64 + threadIdx.x)) >= p ? 1.f / (1.f - p) : 0.f;
    }
    keep[threadIdx.x] = k;
  }
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
  for (int i = threadIdx.x; i < 64 * HP; i += blockDim.x) {
    const int pw = i % HP, c = i / HP;
    const long o = (((long)n * 64 + c) * HP + ph) * HP + pw;
    pooled[o] = sv[pw][c] * keep[c];
    arg[o] = sa[pw][c];
  }
}

// grid (13, N): block ph < 12 writes rows 2ph, 2ph+1 of dy2 (26 pixels wide, zeros at w >= 24); block 12 zeroes rows 24, 25.
__global__ void __launch_bounds__(256) k_pool_bwd(const float* __restrict__ dpooled, const float* __restrict__ pooled, const uint8_t* __restrict__ arg,
                                                  float* __restrict__ dy2, float p, unsigned long long seed, unsigned long long stream,
                                                  const long long* __restrict__ step, int training) {
  __shared__ float sg[HP][65];
  __shared__ uint8_t sa[HP][65];
  griddep_launch();
  griddep_wait();
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

// dz1 [P,32], x [N,28,28] -> dw1[32][9], db1[32].  16 pixel lanes (warps) x 32 channels per CTA, one image per CTA
// iteration, 8 gradient rows in flight per warp.  Every CTA leaves its partial sums in a device-resident table and takes a ticket; the CTA that draws the last
// ticket adds the table up and OVERWRITES dw1 / db1 (no atomics on the outputs, nothing to zero beforehand), then
// resets the ticket counter for the next launch.
constexpr int kC1Blocks = 148, kC1Lanes = 16, kC1Fly = 6;
static __device__ float g_c1w_part[kC1Blocks][320];
static __device__ unsigned g_c1w_ticket;

// Fused data-parallel tail (ar.on): the CTA that finishes the sum also all-reduces (averages) the 320 values with the other ranks
// in the same kernel - one-shot push into a [2 halves][W][320] payload block of every peer's symmetric heap, one signal /
// wait on a reserved (channel, block) pair, local sum.  Same protocol and parity double-buffering as k_allreduce_oneshot in
// coll.cu; it replaces a separate all-reduce launch plus a stream hop at the very end of the backward pass.
struct C1AllReduce {
  int on, chan, blk;
  size_t pay_off;           // heap offset of the payload block (identical on every rank)
  DcpDevCtx c;
};

// Low-latency exchange + optimizer update inside the last CTA (ll.on): the 320 sums go to every rank as 8-byte {value, step}
// words - the flag travels WITH the data (an aligned 8-byte store is single-copy atomic), so there is no fence, no separate
// signal and no kernel boundary between "gradient ready" and "gradient averaged": the CTA stores, polls the W words of each
// element until they carry this step's number, averages, writes dw1 / db1 and applies the optimizer update of conv1's
// parameters on the spot.  Block layout at heap offset `off` on every rank: [2 halves][W source ranks][320] x {f32, u32}.
struct C1LL {
  int on, kind;
  size_t off;
  const long long* step;
  const DcpOptEntry* rows;    // [0] conv1.weight (288 elements), [1] conv1.bias (32)
  const float* hp;
  DcpDevCtx c;
};
// one aligned 64-bit scalar access each way: value in the low word, flag in the high word
__device__ __forceinline__ void st_relaxed_sys_v2(void* p, uint32_t a, uint32_t b) {
  const unsigned long long w = ((unsigned long long)b << 32) | a;
  asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(w) : "memory");
}
__device__ __forceinline__ uint2 ld_relaxed_sys_v2(const void* p) {
  unsigned long long w;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(w) : "l"(p) : "memory");
  return make_uint2((uint32_t)w, (uint32_t)(w >> 32));
}
template <int KIND> __device__ __forceinline__ void c1_update(const DcpOptEntry& e, int i, float g, const float* hp) {
  const OptHyper H = load_hyper<KIND>(hp);
  float p = e.p[i], s1 = e.s1 ? e.s1[i] : 0.f, s2 = e.s2 ? e.s2[i] : 0.f;
  opt_update_h<KIND>(p, g * hp[7], s1, s2, H);
  e.p[i] = p;
  if (e.s1) e.s1[i] = s1;
  if (e.s2) e.s2[i] = s2;
}

// ws_part / ws_ticket: caller-owned partial table ([gridDim.x][320] floats) and ticket word (zero between launches); null = the
// library's own device-resident pair (one per GPU: launches that use it must not overlap).
// dp.on: instead of writing dw1 / db1 the last CTA pushes the 320 sums into its slot of every rank's payload block and
// signals (ops/dp_push.cuh); the consumer kernel averages and writes the gradients.
__global__ void __launch_bounds__(kC1Lanes * 32) k_conv1_wgrad(const float* __restrict__ dz1, const float* __restrict__ x, float* __restrict__ dw1,
                                                               float* __restrict__ db1, long total_pixels, const C1AllReduce ar,
                                                               float* __restrict__ ws_part, unsigned* __restrict__ ws_ticket, const DpPush dp,
                                                               const C1LL ll) {
  __shared__ float red[kC1Lanes][32][11];
  __shared__ float sx[784];
  __shared__ bool last;
  griddep_launch();
  griddep_wait();
  float (*part_tab)[320] = ws_part ? reinterpret_cast<float (*)[320]>(ws_part) : g_c1w_part;
  unsigned* ticket = ws_ticket ? ws_ticket : &g_c1w_ticket;
  const int c = threadIdx.x & 31, lane = threadIdx.x >> 5;       // c: lane in warp, lane: warp index ("pixel lane")
  // main loop: a thread owns 4 channels (16-byte loads, 8 threads cover the 32 channels of a pixel) of every 4th
  // pixel its warp takes; one warp instruction brings 4 pixels, kC1Fly of them are in flight
  const int c4 = c & 7, sub = c >> 3;
  float acc[4][10];
#pragma unroll
  for (int j = 0; j < 4; ++j)
#pragma unroll
    for (int t = 0; t < 10; ++t) acc[j][t] = 0.f;
  const int nimg = (int)(total_pixels / P1);
  for (int n = blockIdx.x; n < nimg; n += gridDim.x) {
    __syncthreads();
    for (int i = threadIdx.x; i < 784; i += blockDim.x) sx[i] = x[(long)n * 784 + i];
    __syncthreads();
    const float4* dz = reinterpret_cast<const float4*>(dz1 + (long)n * P1 * 32) + c4;
    for (int q0 = lane * 4 + sub; q0 < P1; q0 += kC1Lanes * 4 * kC1Fly) {
      float4 g[kC1Fly];
#pragma unroll
      for (int u = 0; u < kC1Fly; ++u) { const int q = q0 + u * kC1Lanes * 4; g[u] = q < P1 ? dz[q * 8] : make_float4(0.f, 0.f, 0.f, 0.f); }
#pragma unroll
      for (int u = 0; u < kC1Fly; ++u) {
        const int q = min(q0 + u * kC1Lanes * 4, P1 - 1), h = q / HW1, wq = q % HW1;
        const float* xi = sx + h * 28 + wq;
        const float gv[4] = {g[u].x, g[u].y, g[u].z, g[u].w};
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
          for (int s = 0; s < 3; ++s) {
            const float xv = xi[r * 28 + s];
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[j][r * 3 + s] = fmaf(gv[j], xv, acc[j][r * 3 + s]);
          }
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[j][9] += gv[j];
      }
    }
  }
  // the 4 pixel sub-lanes of the warp hold sums for the same channels: fold them, lanes 0-7 then own channel 4*c4 + j
#pragma unroll
  for (int j = 0; j < 4; ++j)
#pragma unroll
    for (int t = 0; t < 10; ++t) {
      float v = acc[j][t];
      v += __shfl_xor_sync(0xffffffffu, v, 8);
      v += __shfl_xor_sync(0xffffffffu, v, 16);
      if (sub == 0) red[lane][c4 * 4 + j][t] = v;
    }
  __syncthreads();
  if (threadIdx.x < 320) {
    const int cc = threadIdx.x / 10, t = threadIdx.x % 10;
    float s = 0.f;
#pragma unroll
    for (int l = 0; l < kC1Lanes; ++l) s += red[l][cc][t];
    part_tab[blockIdx.x][threadIdx.x] = s;
    __threadfence();
  }
  __syncthreads();
  if (threadIdx.x == 0) last = atomicAdd(ticket, 1u) == gridDim.x - 1;
  __syncthreads();
  if (!last) return;
  __threadfence();
  // warp w sums the tables of CTAs w, w + 16, ... (10 coalesced loads per table, all independent), then the 16 warp
  // sums meet in shared memory
  float t10[10];
#pragma unroll
  for (int k = 0; k < 10; ++k) t10[k] = 0.f;
#pragma unroll 5
  for (int b = lane; b < (int)gridDim.x; b += kC1Lanes) {
    const float* part = &part_tab[b][c];
#pragma unroll
    for (int k = 0; k < 10; ++k) t10[k] += __ldcg(part + 32 * k);
  }
  float* red2 = &red[0][0][0];                 // [16][320]
#pragma unroll
  for (int k = 0; k < 10; ++k) red2[lane * 320 + c + 32 * k] = t10[k];
  __syncthreads();
  float tot = 0.f;
  if (threadIdx.x < 320) {
    const int cc = threadIdx.x / 10, t = threadIdx.x % 10;
#pragma unroll
    for (int l = 0; l < kC1Lanes; ++l) tot += red2[l * 320 + threadIdx.x];
    if (!ar.on && !dp.on && !ll.on) { if (t < 9) dw1[cc * 9 + t] = tot; else db1[cc] = tot; }
    else if (ar.on) red2[threadIdx.x] = tot;   // keep for the exchange below (row 0 of red2 is this thread's own column)
  }
  if (threadIdx.x == 0) *ticket = 0;
  if (ll.on) {
    __syncthreads();
    if (threadIdx.x < 320) { const int cc = threadIdx.x / 10, t = threadIdx.x % 10; red2[t < 9 ? cc * 9 + t : 288 + cc] = tot; }   // bucket order
    __syncthreads();
    if (threadIdx.x >= 320) return;
    const DcpDevCtx& cx = ll.c;
    const int W = cx.world, i = threadIdx.x;
    const uint32_t epoch = (uint32_t)(*ll.step);
    const size_t half_off = ll.off + (size_t)(epoch & 1u) * W * 320 * 8;
    const float mine = red2[i];
    for (int p = 0; p < W; ++p) st_relaxed_sys_v2(cx.peer[p] + half_off + ((size_t)cx.rank * 320 + i) * 8, __float_as_uint(mine), epoch);
    const char* slots = cx.peer[cx.rank] + half_off + (size_t)i * 8;
    float sum = 0.f;
    bool ok = true;
    const unsigned long long tw0 = i == 0 ? globaltimer_ns() : 0ull;
    for (int p = 0; p < W && ok; ++p) {
      const char* w = slots + (size_t)p * 320 * 8;
      uint2 v = ld_relaxed_sys_v2(w);
      if (v.y != epoch) {
        const unsigned long long t0 = globaltimer_ns();
        for (;;) {
#pragma unroll 1
          for (int k = 0; k < 64 && v.y != epoch; ++k) v = ld_relaxed_sys_v2(w);
          if (v.y == epoch) break;
          if (globaltimer_ns() - t0 > cx.timeout_ns) { if (cx.err) *((volatile int*)cx.err) = 1; ok = false; break; }
        }
      }
      sum += __uint_as_float(v.x);
    }
    if (!ok) return;          // a peer never delivered: leave the parameters alone, the host sees the error flag
    // bookkeeping for bench/step_ablation.py: total time element 0 spent waiting for the peers' words (= rank skew + NVLink flight)
    if (i == 0) *reinterpret_cast<unsigned long long*>(ticket + 2) += globaltimer_ns() - tw0;
    sum *= 1.f / (float)W;
    const DcpOptEntry e = ll.rows[i < 288 ? 0 : 1];
    const int j = i < 288 ? i : i - 288;
    const_cast<float*>(static_cast<const float*>(e.g))[j] = sum;      // .grad holds the averaged gradient, as after DDP's all-reduce
    if (ll.kind == 0) c1_update<0>(e, j, sum, ll.hp); else if (ll.kind == 1) c1_update<1>(e, j, sum, ll.hp); else c1_update<2>(e, j, sum, ll.hp);
    return;
  }
  if (dp.on) {
    // push-only tail: the sums go out in the order the gradient bucket stores them (dw1 [32][9], then db1 [32])
    __syncthreads();
    if (threadIdx.x < 320) { const int cc = threadIdx.x / 10, t = threadIdx.x % 10; red2[t < 9 ? cc * 9 + t : 288 + cc] = tot; }
    __syncthreads();
    const size_t slot_off = dp_slot_off(dp, (unsigned)(*dp.step) & 1u);
    if (threadIdx.x < 80) dp_push4(dp, slot_off, (size_t)threadIdx.x * 4, *reinterpret_cast<const float4*>(red2 + threadIdx.x * 4));
    __syncthreads();
    dp_signal(dp);
    return;
  }
  if (!ar.on) return;
  // ---- all-reduce (AVG) of the 320 sums across the ranks, by this one CTA ----
  const DcpDevCtx& cx = ar.c;
  const int W = cx.world;
  char* me = cx.peer[cx.rank];
  __shared__ uint32_t s_half, s_epoch;
  if (threadIdx.x == 0) { s_half = *((volatile uint32_t*)dcp_calls(me, ar.chan, ar.blk)) & 1u; s_epoch = *((volatile uint32_t*)dcp_epoch(me, ar.chan, ar.blk)) + 1; }
  __syncthreads();
  const size_t half_off = ar.pay_off + (size_t)s_half * W * 320 * sizeof(float);
  if (threadIdx.x < 320) {
    const float v = red2[threadIdx.x];
    for (int p = 0; p < W; ++p) reinterpret_cast<float*>(cx.peer[p] + half_off)[cx.rank * 320 + threadIdx.x] = v;
  }
  __syncthreads();
  int arrived = 1;
  if (threadIdx.x < W) {
    st_release_sys(dcp_sig(cx.peer[threadIdx.x], ar.chan, ar.blk, cx.rank), s_epoch);
    arrived = wait_ge(cx, dcp_sig(me, ar.chan, ar.blk, threadIdx.x), s_epoch) ? 1 : 0;
  }
  if (!__syncthreads_and(arrived)) return;      // a peer never delivered: leave the gradient un-reduced, the error flag is raised
  if (threadIdx.x < 320) {
    const float* slots = reinterpret_cast<const float*>(me + half_off);
    float sum = 0.f;
    for (int p = 0; p < W; ++p) sum += __ldcg(slots + p * 320 + threadIdx.x);
    sum *= 1.f / (float)W;
    const int cc = threadIdx.x / 10, t = threadIdx.x % 10;
    if (t < 9) dw1[cc * 9 + t] = sum; else db1[cc] = sum;
  }
  if (threadIdx.x == 0) {
    *((volatile uint32_t*)dcp_epoch(me, ar.chan, ar.blk)) = s_epoch;
    volatile uint32_t* calls = dcp_calls(me, ar.chan, ar.blk); *calls = *calls + 1;
  }
}

}  // namespace

extern "C" {

int dcp_convnet_conv1_fwd(const float* x, const float* w, const float* b, float* a1, int N, cudaStream_t s) {
  const long pixels = (long)N * P1;
  DCP_CUDA_CHECK(dcp_launch_pdl(k_conv1_fwd, dim3((unsigned)((pixels * 2 + 255) / 256)), dim3(256), 0, s, x, w, b, a1, pixels)); return 0;
}
int dcp_convnet_pool_fwd(const float* a2, float* pooled, void* arg, int N, float p, unsigned long long seed, unsigned long long stream,
                         const void* step, int training, cudaStream_t s) {
  DCP_CUDA_CHECK(dcp_launch_pdl(k_pool_fwd, dim3(HP, N), dim3(256), 0, s, a2, pooled, (uint8_t*)arg, p, seed, stream, (const long long*)step, training)); return 0;
}
int dcp_convnet_pool_bwd(const float* dpooled, const float* pooled, const void* arg, float* dy2, int N, float p, unsigned long long seed,
                         unsigned long long stream, const void* step, int training, cudaStream_t s) {
  DCP_CUDA_CHECK(dcp_launch_pdl(k_pool_bwd, dim3(HP + 1, N), dim3(256), 0, s, dpooled, pooled, (const uint8_t*)arg, dy2, p, seed, stream, (const long long*)step, training)); return 0;
}
// dw1 / db1 are overwritten.  The partial-sum table is one per device: two of these launches running at the same time
// (different streams) would race on it.
int dcp_convnet_conv1_wgrad(const float* dz1, const float* x, float* dw1, float* db1, int N, cudaStream_t s) {
  C1AllReduce ar{};
  DpPush dp{};
  DCP_CUDA_CHECK(dcp_launch_pdl(k_conv1_wgrad, dim3(N < kC1Blocks ? N : kC1Blocks), dim3(kC1Lanes * 32), 0, s, dz1, x, dw1, db1, (long)N * P1, ar,
                                (float*)nullptr, (unsigned*)nullptr, dp, C1LL{})); return 0;
}
// The same kernel with the data-parallel all-reduce (AVG over the ranks of symmetric heap `h`) of dw1 / db1 fused into its last
// CTA.  payload: a block of 2 * world * 320 floats INSIDE the symmetric heap, allocated in the same order on every rank and
// used by nothing else; (chan, blk): a signal-pad slot no other kernel uses (the trainer passes CH_DDP, DCP_MAX_BLOCKS - 1).
// Every rank must make this call the same number of times.  dw1 / db1 receive the averaged gradients.
const DcpDevCtx* dcp_symm_ctx(void* h);
size_t dcp_symm_bytes(void* h);
int dcp_convnet_conv1_wgrad_allreduce(void* h, const float* dz1, const float* x, float* dw1, float* db1, int N, void* payload, int chan, int blk,
                                      cudaStream_t s) {
  const DcpDevCtx& c = *dcp_symm_ctx(h);
  const char* base = c.peer[c.rank];
  const char* pay = static_cast<const char*>(payload);
  if (pay < base || pay + (size_t)2 * c.world * 320 * 4 > base + dcp_symm_bytes(h) || ((pay - base) & 15)) {
    dcp_set_error("conv1 wgrad + all-reduce: the payload block must be a 16-byte aligned region of the symmetric heap"); return -1;
  }
  if (chan < 0 || chan >= DCP_NUM_CHANNELS || blk < 0 || blk >= DCP_MAX_BLOCKS) { dcp_set_error("conv1 wgrad + all-reduce: bad signal slot (%d, %d)", chan, blk); return -1; }
  C1AllReduce ar{};
  ar.on = c.world > 1; ar.chan = chan; ar.blk = blk; ar.pay_off = (size_t)(pay - base); ar.c = c;
  DpPush dp{};
  DCP_CUDA_CHECK(dcp_launch_pdl(k_conv1_wgrad, dim3(N < kC1Blocks ? N : kC1Blocks), dim3(kC1Lanes * 32), 0, s, dz1, x, dw1, db1, (long)N * P1, ar,
                                (float*)nullptr, (unsigned*)nullptr, dp, C1LL{})); return 0;
}
// Push-only variant for the fused data-parallel tail (ops/dp_push.cuh, convnet_dp.cu): the 320 sums go into this rank's slot
// (floats [elem_off, elem_off + 320) in bucket order: dw1 then db1) of every rank's payload block, followed by one signal per
// rank on (chan, blk); dw1 / db1 are NOT written - dcp_dp_tail produces the averaged gradients.  ws: caller-owned scratch of
// 4 + 148 * 320 floats (ticket word + per-CTA partial sums), zero before the first call; one per concurrent user.
int dcp_convnet_conv1_wgrad_push(void* h, const float* dz1, const float* x, int N, void* payload, unsigned slot_elems, unsigned elem_off,
                                 int chan, int blk, const void* step, void* ws, cudaStream_t s) {
  const DcpDevCtx& c = *dcp_symm_ctx(h);
  const char* base = c.peer[c.rank];
  const char* pay = static_cast<const char*>(payload);
  if (pay < base || pay + (size_t)2 * c.world * slot_elems * 4 > base + dcp_symm_bytes(h) || ((pay - base) & 15) || (slot_elems & 3) || (elem_off & 3) ||
      elem_off + 320 > slot_elems) {
    dcp_set_error("conv1 wgrad push: the payload block must be a 16-byte aligned region of the symmetric heap holding 2 * world * slot_elems floats"); return -1;
  }
  if (chan < 0 || chan >= DCP_NUM_CHANNELS || blk < 0 || blk >= DCP_MAX_BLOCKS || !ws || !step) { dcp_set_error("conv1 wgrad push: bad arguments"); return -1; }
  C1AllReduce ar{};
  DpPush dp{};
  dp.on = 1; dp.chan = chan; dp.blk = blk; dp.nvls = dp_use_nvls(c); dp.pay_off = (size_t)(pay - base); dp.slot_elems = slot_elems; dp.elem_off = elem_off;
  dp.step = static_cast<const long long*>(step); dp.ticket = nullptr; dp.c = c;
  unsigned* ticket = static_cast<unsigned*>(ws);
  float* part = static_cast<float*>(ws) + 4;
  DCP_CUDA_CHECK(dcp_launch_pdl(k_conv1_wgrad, dim3(N < kC1Blocks ? N : kC1Blocks), dim3(kC1Lanes * 32), 0, s, dz1, x, (float*)nullptr, (float*)nullptr,
                                (long)N * P1, ar, part, ticket, dp, C1LL{})); return 0;
}
// conv1's weight-gradient, its data-parallel average and its optimizer update in ONE kernel (struct C1LL above).  ll_block: a
// region of the symmetric heap holding 2 * world * 320 * 8 bytes, at the same offset on every rank, zeroed before first use and
// used by nothing else; rows: two rows of the optimizer's device pointer table - conv1.weight, conv1.bias - whose grad pointers
// receive the averaged gradient; hp / kind as for dcp_optim_step (dcp_optim_advance'd by the caller); step: device int64
// counter holding the same (non-zero, increasing) value on every rank when the kernel runs; ws: 4 + 148 * 320 floats of
// caller-owned scratch, zero before the first call.
int dcp_convnet_conv1_wgrad_ll(void* h, const float* dz1, const float* x, int N, void* ll_block, const void* rows, const float* hp, int kind,
                               const void* step, void* ws, cudaStream_t s) {
  const DcpDevCtx& c = *dcp_symm_ctx(h);
  const char* base = c.peer[c.rank];
  const char* blk = static_cast<const char*>(ll_block);
  if (blk < base || blk + (size_t)2 * c.world * 320 * 8 > base + dcp_symm_bytes(h) || ((blk - base) & 15)) {
    dcp_set_error("conv1 wgrad ll: ll_block must be a 16-byte aligned region of the symmetric heap holding 2 * world * 320 * 8 bytes"); return -1;
  }
  if (!rows || !hp || !step || !ws || kind < 0 || kind > 2) { dcp_set_error("conv1 wgrad ll: bad arguments"); return -1; }
  C1LL ll{};
  ll.on = 1; ll.kind = kind; ll.off = (size_t)(blk - base); ll.step = static_cast<const long long*>(step);
  ll.rows = static_cast<const DcpOptEntry*>(rows); ll.hp = hp; ll.c = c;
  unsigned* ticket = static_cast<unsigned*>(ws);
  float* part = static_cast<float*>(ws) + 4;
  DCP_CUDA_CHECK(dcp_launch_pdl(k_conv1_wgrad, dim3(N < kC1Blocks ? N : kC1Blocks), dim3(kC1Lanes * 32), 0, s, dz1, x, (float*)nullptr, (float*)nullptr,
                                (long)N * P1, C1AllReduce{}, part, ticket, DpPush{}, ll)); return 0;
}

}  // extern "C"

// =====================================================================================================
// Head of the network in ONE kernel, forward AND backward (K9-K14 and their gradients, ref:main.py:40-44,61):
//   h1[N,128] -> BatchNorm1d (batch statistics, running-stat update) -> ReLU -> dropout p2 -> fc2 (128 -> 10)
//   -> log_softmax + nll (mean) -> d logits -> fc2 weight/bias/data gradients -> dropout / ReLU backward
//   -> BatchNorm backward -> dh1[N,128], dgamma, dbeta.
// The whole problem is 16 K activations.  BatchNorm, ReLU and dropout are independent per feature, only fc2 mixes
// features: ONE thread-block cluster of 8 CTAs splits the 128 features 16 per CTA, each CTA computes its partial
// logits, and the partials meet through distributed shared memory (every CTA reads the other seven, then finishes the
// softmax redundantly so no second exchange is needed).  Two cluster barriers in total.  N <= 128 rows.
// Thread (fl = tid % 16, rg = tid / 16) owns feature f0 + fl of rows rg, rg + 32, rg + 64, rg + 96.
// =====================================================================================================
#include <cooperative_groups.h>
namespace {
namespace cg = cooperative_groups;
constexpr int HF = 128, HC = 10, HMAXN = 128;
constexpr int HCL = 8, HFL = HF / HCL;          // cluster size, features per CTA
constexpr int HT = 512, HRG = HT / HFL, HR = HMAXN / HRG;   // threads, row groups (32), rows per thread (4)

// Data-parallel BatchNorm buffers: `off` = heap offset of a block [running_mean 128][running_var 128][num_batches_tracked int64]
// that exists at the same offset on every rank; run_mean / run_var / nbt of the launch point into this rank's copy.
struct HeadBnPush { int on; size_t off; DcpDevCtx c; };

struct __align__(16) HeadSmem {
  float xhat[HMAXN][HFL + 1];
  float h3[HMAXN][HFL + 1];
  float w2[HC][HFL];
  float part[HMAXN][HC];          // this CTA's partial logits (read by the whole cluster); 16-byte aligned offset
  float logit[HMAXN][HC];         // sum of the 8 partial logits
  float w2t[HFL][HC + 1];         // w2 slice transposed (conflict-free in the partial-logit loop)
  float dlog[HMAXN][HC + 2];
  float red[HRG / 2][HFL][HC + 1];
  float stat[4][HFL];             // mean, rstd, dbeta/N, dgamma/N
  float loss;
};

// sum over the 32 row groups of a per-thread value: the two row groups of a warp by shuffle, the 16 warps through smem
__device__ __forceinline__ float head_colsum(HeadSmem& S, float v, int slot, int fl, int warp) {
  v += __shfl_xor_sync(0xffffffffu, v, 16);
  S.red[warp][fl][slot] = v;
  return v;
}


__global__ void __cluster_dims__(HCL, 1, 1) __launch_bounds__(HT, 1)
k_convnet_head(const float* __restrict__ h1, const long long* __restrict__ target, const float* __restrict__ gamma, const float* __restrict__ beta,
               float* __restrict__ run_mean, float* __restrict__ run_var, long long* __restrict__ nbt, const float* __restrict__ w2,
               const float* __restrict__ b2, float* __restrict__ loss_out, float* __restrict__ logits_out, float* __restrict__ dh1,
               float* __restrict__ dgamma, float* __restrict__ dbeta, float* __restrict__ dw2, float* __restrict__ db2,
               float* __restrict__ dbias1, int N, float eps,
               float momentum, float p2, unsigned long long seed, unsigned long long stream, const long long* __restrict__ step, int training,
               const HeadBnPush bp) {
  __shared__ HeadSmem S;
  griddep_launch();
  griddep_wait();
  cg::cluster_group cluster = cg::this_cluster();
  const int rank = (int)cluster.block_rank();
  const int tid = threadIdx.x, fl = tid & (HFL - 1), rg = tid >> 4, warp = tid >> 5, lane = tid & 31;
  const int f0 = rank * HFL, f = f0 + fl;
  const float invN = 1.f / (float)N;
  float x[HR];
  // ---- 1. batch statistics (two-pass, data stays in registers) ----
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < HR; ++i) { const int r = rg + i * HRG; x[i] = r < N ? h1[r * HF + f] : 0.f; s += x[i]; }
  // every other global operand is fetched now, so that no later phase starts with an exposed memory round trip
  const float gm = gamma[f], bt = beta[f];
  const unsigned long long st = stream + (step ? (unsigned long long)(*step) * 0x10000ull : 0ull);
  const int tgt = (tid >> 2) < N ? (int)target[tid >> 2] : -1;
  float bias2[HC];
#pragma unroll
  for (int j = 0; j < HC; ++j) bias2[j] = b2[j];
  head_colsum(S, s, 0, fl, warp);
  if (tid < HC * HFL) { const float wv = w2[(tid / HFL) * HF + f0 + tid % HFL]; S.w2[tid / HFL][tid % HFL] = wv; S.w2t[tid % HFL][tid / HFL] = wv; }
  if (tid == 0) S.loss = 0.f;
  __syncthreads();
  if (tid < HFL) { float t = 0.f; for (int g = 0; g < HRG / 2; ++g) t += S.red[g][tid][0]; S.stat[0][tid] = t * invN; }
  __syncthreads();
  const float mean = S.stat[0][fl];
  float v = 0.f;
#pragma unroll
  for (int i = 0; i < HR; ++i) { const int r = rg + i * HRG; if (r < N) { const float d = x[i] - mean; v += d * d; } }
  head_colsum(S, v, 1, fl, warp);
  __syncthreads();
  if (tid < HFL) {
    float t = 0.f; for (int g = 0; g < HRG / 2; ++g) t += S.red[g][tid][1];
    const float var = t * invN, mu = S.stat[0][tid];
    S.stat[1][tid] = rsqrtf(var + eps);
    if (training && run_mean) {
      const float unbiased = N > 1 ? var * (float)N / (float)(N - 1) : var;
      const float nm = (1.f - momentum) * run_mean[f0 + tid] + momentum * mu;
      const float nv = (1.f - momentum) * run_var[f0 + tid] + momentum * unbiased;
      if (!bp.on) { run_mean[f0 + tid] = nm; run_var[f0 + tid] = nv; }
      else if (bp.c.rank == 0) {
        // data-parallel buffer sync folded into the kernel: rank 0's running statistics are THE running statistics
        // (DistributedDataParallel broadcasts rank 0's buffers before every forward); rank 0 writes them into every
        // rank's copy, the other ranks leave theirs alone
        for (int p = 0; p < bp.c.world; ++p) {
          float* dst = reinterpret_cast<float*>(bp.c.peer[p] + bp.off);
          dst[f0 + tid] = nm; dst[HF + f0 + tid] = nv;
        }
      }
    }
  }
  if (rank == 0 && tid == 0 && training && nbt) {
    if (!bp.on) *nbt += 1;
    else if (bp.c.rank == 0) {
      const long long nb = *nbt + 1;
      for (int p = 0; p < bp.c.world; ++p) *reinterpret_cast<long long*>(bp.c.peer[p] + bp.off + 2 * HF * sizeof(float)) = nb;
    }
  }
  __syncthreads();
  // ---- 2. normalise + affine + ReLU + dropout ----
  const float rstd = S.stat[1][fl];
  const bool drop = training && p2 > 0.f;
  const float dscale = drop ? 1.f / (1.f - p2) : 1.f;
  // one Philox call yields the uniforms of 4 consecutive features of a row: lane q of a feature quad generates row q of
  // its four rows and the quad trades the words by shuffle (same element -> same number as philox_uniform(idx))
  uint4 rnd = make_uint4(0, 0, 0, 0);
  if (drop) { Philox ph(seed); rnd = ph((unsigned long long)(((rg + (fl & 3) * HRG) * HF + f) >> 2), st); }
  unsigned keep = 0, pos = 0;   // bit i: dropout kept / ReLU active for row rg + 32 i
#pragma unroll
  for (int i = 0; i < HR; ++i) {
    const int r = rg + i * HRG;
    const int src = (lane & ~3) | i;
    const uint32_t wx = __shfl_sync(0xffffffffu, rnd.x, src), wy = __shfl_sync(0xffffffffu, rnd.y, src);
    const uint32_t wz = __shfl_sync(0xffffffffu, rnd.z, src), ww = __shfl_sync(0xffffffffu, rnd.w, src);
    const uint32_t wsel = (fl & 3) == 0 ? wx : (fl & 3) == 1 ? wy : (fl & 3) == 2 ? wz : ww;
    float xh = 0.f, h = 0.f;
    if (r < N) {
      xh = (x[i] - mean) * rstd;
      float a = xh * gm + bt;
      if (a > 0.f) pos |= 1u << i; else a = 0.f;
      const bool k = !drop || (float)(wsel >> 8) * (1.0f / 16777216.0f) >= p2;
      if (k) keep |= 1u << i;
      h = k ? a * dscale : 0.f;
    }
    S.xhat[r][fl] = xh;
    S.h3[r][fl] = h;
  }
  __syncthreads();
  // ---- 3a. partial logits over this CTA's 16 features: thread (row r, q) computes classes q, q + 4, q + 8 ----
  {
    const int r = tid >> 2, q = tid & 3;
    float a0 = 0.f, a1 = 0.f, a2 = 0.f;
#pragma unroll
    for (int k = 0; k < HFL; ++k) {
      const float hv = S.h3[r][k];
      a0 = fmaf(hv, S.w2t[k][q], a0);
      a1 = fmaf(hv, S.w2t[k][q + 4], a1);
      if (q < 2) a2 = fmaf(hv, S.w2t[k][q + 8], a2);
    }
    S.part[r][q] = a0; S.part[r][q + 4] = a1;
    if (q < 2) S.part[r][q + 8] = a2;
  }
  cluster.sync();
  // ---- 3b. every CTA adds up the 8 partial-logit tables with coalesced 16-byte DSMEM loads (one table = 5 KB
  // contiguous), then 4 threads per row finish log-softmax + nll + dlogits from shared memory ----
  if (tid < HMAXN * HC / 4) {
    float4 acc4 = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int c = 0; c < HCL; ++c) {
      const float4 t = reinterpret_cast<const float4*>(cluster.map_shared_rank(&S.part[0][0], c))[tid];
      acc4.x += t.x; acc4.y += t.y; acc4.z += t.z; acc4.w += t.w;
    }
    reinterpret_cast<float4*>(&S.logit[0][0])[tid] = acc4;
  }
  __syncthreads();
  {
    const int r = tid >> 2, q = tid & 3;
    float acc[HC];
#pragma unroll
    for (int j = 0; j < HC; ++j) acc[j] = S.logit[r][j] + bias2[j];
    float m = acc[0];
#pragma unroll
    for (int j = 1; j < HC; ++j) m = fmaxf(m, acc[j]);
    float se = 0.f;
#pragma unroll
    for (int j = 0; j < HC; ++j) se += __expf(acc[j] - m);
    const float lse = m + __logf(se);
    const int t = tgt;
    // thread q of the row writes columns q, q + 4, q + 8
#pragma unroll
    for (int j = 0; j < HC; ++j) {
      if ((j & 3) == q) {
        const float pj = __expf(acc[j] - lse);
        S.dlog[r][j] = r < N ? (pj - (j == t ? 1.f : 0.f)) * invN : 0.f;
        if (rank == 0 && logits_out && r < N) logits_out[r * HC + j] = acc[j] - lse;   // log-probabilities, as the model's forward returns
      }
    }
    if (rank == 0) {                       // mean nll: one shared-memory atomic per warp
      float at = acc[0];
#pragma unroll
      for (int j = 1; j < HC; ++j) at = j == t ? acc[j] : at;
      const float lv = warp_sum((q == 0 && t >= 0 && t < HC) ? (lse - at) * invN : 0.f);
      if (lane == 0) atomicAdd(&S.loss, lv);
    }
  }
  __syncthreads();
  if (rank == 0 && tid == 0) *loss_out = S.loss;
  // ---- 4. fc2 gradients (this CTA's 16 columns of dw2; db2 on CTA 0) ----
  float gw[HC];
#pragma unroll
  for (int j = 0; j < HC; ++j) gw[j] = 0.f;
#pragma unroll
  for (int i = 0; i < HR; ++i) {
    const int r = rg + i * HRG;
    const float hv = S.h3[r][fl];                     // rows >= N hold zeros
#pragma unroll
    for (int j = 0; j < HC; ++j) gw[j] = fmaf(S.dlog[r][j], hv, gw[j]);
  }
#pragma unroll
  for (int j = 0; j < HC; ++j) head_colsum(S, gw[j], j, fl, warp);
  // ---- 5. dh3 -> dropout / ReLU backward -> BatchNorm backward ----
  float g[HR];
  float sb = 0.f, sg = 0.f;
#pragma unroll
  for (int i = 0; i < HR; ++i) {
    const int r = rg + i * HRG;
    float gv = 0.f;
    if (r < N && ((keep >> i) & 1u) && ((pos >> i) & 1u)) {
#pragma unroll
      for (int j = 0; j < HC; ++j) gv = fmaf(S.dlog[r][j], S.w2[j][fl], gv);
      gv *= dscale;
    }
    g[i] = gv;
    sb += gv;
    sg += gv * S.xhat[r][fl];
  }
  __syncthreads();
  if (tid < HC * HFL) {
    const int j = tid / HFL, ff = tid % HFL;
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < HRG / 2; ++w) t += S.red[w][ff][j];
    dw2[j * HF + f0 + ff] = t;
  }
  if (rank == 0 && warp >= 16 - HC) {    // db2: warp per class, 4 rows per lane (rows >= N hold zeros)
    const int j = warp - (16 - HC);
    const float t = warp_sum(S.dlog[lane][j] + S.dlog[lane + 32][j] + S.dlog[lane + 64][j] + S.dlog[lane + 96][j]);
    if (lane == 0) db2[j] = t;
  }
  __syncthreads();                       // S.red is reused
  head_colsum(S, sb, 0, fl, warp);
  head_colsum(S, sg, 1, fl, warp);
  __syncthreads();
  if (tid < HFL) {
    float tb = 0.f, tg = 0.f;
    for (int w = 0; w < HRG / 2; ++w) { tb += S.red[w][tid][0]; tg += S.red[w][tid][1]; }
    dbeta[f0 + tid] = tb; dgamma[f0 + tid] = tg;
    S.stat[2][tid] = tb * invN; S.stat[3][tid] = tg * invN;
  }
  __syncthreads();
  const float k1 = gm * rstd, k2 = S.stat[2][fl], k3 = S.stat[3][fl];
  float sdh = 0.f;
#pragma unroll
  for (int i = 0; i < HR; ++i) {
    const int r = rg + i * HRG;
    if (r < N) { const float d = k1 * (g[i] - k2 - S.xhat[r][fl] * k3); dh1[r * HF + f] = d; sdh += d; }
  }
  if (dbias1) {                          // gradient of the bias of the layer that produced h1 (column sums of dh1)
    head_colsum(S, sdh, 2, fl, warp);
    __syncthreads();
    if (tid < HFL) { float t = 0.f; for (int w = 0; w < HRG / 2; ++w) t += S.red[w][tid][2]; dbias1[f0 + tid] = t; }
  }
  cluster.sync();                        // nobody leaves while a peer may still read its partial logits
}
}  // namespace

extern "C" {
// Fused head (see above).  N <= 128.  logits_out (log-probabilities [N,10]) may be null; dbias1 ([128], may be null)
// receives the column sums of dh1, i.e. the bias gradient of the Linear layer in front.  All gradient outputs are
// fully overwritten.  Training mode only (batch statistics); p2 = dropout probability of the second dropout.
int dcp_convnet_head(const float* h1, const void* target, const float* gamma, const float* beta, float* run_mean, float* run_var, void* nbt,
                     const float* w2, const float* b2, float* loss, float* logits_out, float* dh1, float* dgamma, float* dbeta, float* dw2,
                     float* db2, float* dbias1, int N, float eps, float momentum, float p2, unsigned long long seed,
                     unsigned long long stream_id, const void* step, int training, cudaStream_t s) {
  if (N < 1 || N > HMAXN) { dcp_set_error("convnet head: batch %d not in [1, %d]", N, HMAXN); return -1; }
  HeadBnPush bp{};
  DCP_CUDA_CHECK(dcp_launch_pdl(k_convnet_head, dim3(HCL), dim3(HT), 0, s, h1, (const long long*)target, gamma, beta, run_mean, run_var,
                                (long long*)nbt, w2, b2, loss, logits_out, dh1, dgamma, dbeta, dw2, db2, dbias1, N, eps, momentum, p2, seed,
                                stream_id, (const long long*)step, training, bp));
  return 0;
}
// The same kernel inside a data-parallel job: `bn_block` is a 1032-byte block of symmetric heap `h` laid out
// [running_mean 128 f32][running_var 128 f32][num_batches_tracked i64], allocated at the same offset on every rank.  Rank 0
// computes the new running statistics from its batch and writes them into EVERY rank's block (peer stores over NVLink from
// inside the kernel); the other ranks do not touch theirs.  That is what DistributedDataParallel's per-forward buffer
// broadcast from rank 0 amounts to - without a collective launch, a side stream or a join in front of the head.
// A rank may read the block after any later cross-rank synchronisation of the step (the gradient exchange is one).
int dcp_convnet_head_dp(void* h, const float* h1, const void* target, const float* gamma, const float* beta, void* bn_block,
                        const float* w2, const float* b2, float* loss, float* logits_out, float* dh1, float* dgamma, float* dbeta, float* dw2,
                        float* db2, float* dbias1, int N, float eps, float momentum, float p2, unsigned long long seed,
                        unsigned long long stream_id, const void* step, int training, cudaStream_t s) {
  if (N < 1 || N > HMAXN) { dcp_set_error("convnet head: batch %d not in [1, %d]", N, HMAXN); return -1; }
  const DcpDevCtx& c = *dcp_symm_ctx(h);
  char* base = c.peer[c.rank];
  char* blk = static_cast<char*>(bn_block);
  if (blk < base || blk + 2 * HF * 4 + 8 > base + dcp_symm_bytes(h) || ((blk - base) & 15)) {
    dcp_set_error("convnet head (data-parallel): bn_block must be a 16-byte aligned region of the symmetric heap"); return -1;
  }
  HeadBnPush bp{};
  bp.on = c.world > 1; bp.off = (size_t)(blk - base); bp.c = c;
  float* rm = reinterpret_cast<float*>(blk);
  DCP_CUDA_CHECK(dcp_launch_pdl(k_convnet_head, dim3(HCL), dim3(HT), 0, s, h1, (const long long*)target, gamma, beta, rm, rm + HF,
                                reinterpret_cast<long long*>(blk + 2 * HF * 4), w2, b2, loss, logits_out, dh1, dgamma, dbeta, dw2, db2, dbias1, N,
                                eps, momentum, p2, seed, stream_id, (const long long*)step, training, bp));
  return 0;
}
}  // extern "C"
