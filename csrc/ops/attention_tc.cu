// Causal flash attention on tcgen05 / TMEM / TMA for the GPT-2 blocks (bf16, head dim 64): forward and backward.
// (Not in the reference - it has no attention; BASELINE.json configs[4] trains GPT-2 small, whose stock PyTorch arm runs
// F.scaled_dot_product_attention, i.e. cuDNN's flash kernels.  This file is what replaces that library call.)
//
// Inputs are read IN PLACE from the fused qkv projection [B*T, 3*C] (C = H*64): q / k / v of head h are the 64-column slices
// at column h*64, C + h*64, 2C + h*64 - one 2-D TMA tensor map with a {64, 128} box serves all three, no split / transpose
// copies.  The output is written as [B*T, C] (heads merged), which is what the projection GEMM consumes.
//
// Forward, one CTA per (128-row query block i, batch*head), key/value blocks j = 0..i:
//   warp 4   TMA producer: Q once, then K_0, V_0, K_1, V_1, ... through a ring of three 16 KB tiles
//   warp 5   MMA issuer:   S_j = Q K_j^T (128x128x64) into TMEM buffer j & 1; PV_j = P_j V_j (128x64x128) into the first 64
//                          columns of the SAME buffer once the softmax has consumed S_j (256 TMEM columns and ~98 KB of shared
//                          memory per CTA: two CTAs per SM, which is where the MMA / softmax overlap comes from)
//   warps 0-3 softmax:     thread = query row (TMEM lane).  Online softmax in the log2 domain: pass 1 reads S_j for the row
//                          maximum, pass 2 reads it again to produce P_j (bf16) straight into the 128-byte-swizzled K-major
//                          shared-memory tile the PV MMA reads.  The output row lives in REGISTERS (64 floats): each PV_j is
//                          added from TMEM and rescaled there, so TMEM never has to be rewritten.
// The row-wise log-sum-exp (log2 domain) is saved for the backward pass.
#include <cudaTypedefs.h>

#include "ops/tc_common.cuh"

namespace {
using namespace tc;

constexpr int kD = 64, kBQ = 128, kBK = 128;
constexpr int kAttnThreads = 192;
constexpr int kTileBytes = kBQ * kD * 2;            // 16 KB: one 128 x 64 bf16 tile (rows of 128 B, 128-byte swizzle)
constexpr int kRing = 3;                            // K / V tiles in flight

struct __align__(16) AttnBars {
  uint64_t q_full, kv_full[kRing], kv_empty[kRing], s_full[2], p_ready, pv_full;
  uint32_t tmem_base;
};

struct AttnFwdParams {
  int T, H, C;                 // sequence length, heads, model width (H * 64)
  float scale_log2;            // softmax scale * log2(e)
  __nv_bfloat16* out;          // [B*T, C]
  float* lse;                  // [B*H, T] log2-domain log-sum-exp
};

// 16-byte unit `u` (0..7) of row r inside a 128-row x 128-byte swizzled chunk
__device__ __forceinline__ uint32_t swz(uint32_t r, uint32_t u) { return r * 128u + ((u ^ (r & 7u)) << 4); }

__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ float ex2(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}

__global__ void __launch_bounds__(kAttnThreads, 2)
k_attn_fwd(const __grid_constant__ CUtensorMap map_qkv, const AttnFwdParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = align_smem_1024(smem_raw);
  uint8_t* q_smem = smem;                                   // 16 KB
  uint8_t* kv_smem = smem + kTileBytes;                     // kRing x 16 KB: K_0, V_0, K_1, V_1, ...
  uint8_t* p_smem = kv_smem + kRing * kTileBytes;           // 32 KB: P as two 64-column K-major chunks
  AttnBars* bars = reinterpret_cast<AttnBars*>(p_smem + 2 * kTileBytes);
  const int warp = warp_idx_uniform(), lane = threadIdx.x & 31;
  const int qi = (int)gridDim.x - 1 - (int)blockIdx.x;      // heavy (late) query blocks first
  const int bh = blockIdx.y, b = bh / p.H, h = bh % p.H;
  const int row0 = b * p.T + qi * kBQ;                      // first row of this query block in the [B*T, 3C] matrix
  const int nkv = qi + 1;

  griddep_launch();
  if (warp == 4 && lane == 0) tma_prefetch_desc(&map_qkv);
  if (warp == 5) {
    if (lane == 0) {
      mbar_init(&bars->q_full, 1);
      for (int i = 0; i < kRing; ++i) { mbar_init(&bars->kv_full[i], 1); mbar_init(&bars->kv_empty[i], 1); }
      for (int i = 0; i < 2; ++i) mbar_init(&bars->s_full[i], 1);
      mbar_init(&bars->p_ready, 128); mbar_init(&bars->pv_full, 1);
      mbar_fence_init();
    }
    __syncwarp();
    tmem_alloc<256>(&bars->tmem_base);
  }
  griddep_wait();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *((volatile uint32_t*)&bars->tmem_base);
  const uint32_t tS0 = tmem;                                // S_j (128 columns) and later P_j V_j (64 columns) in buffer j & 1

  if (warp == 4) {
    // ---- TMA producer
    if (lane == 0) {
      mbar_expect_tx(&bars->q_full, kTileBytes);
      tma_load_2d(q_smem, &map_qkv, &bars->q_full, h * kD, row0);
    }
    __syncwarp();
    for (int t = 0; t < 2 * nkv; ++t) {                     // tile t: K_j for t = 2 j, V_j for t = 2 j + 1
      const int st = t % kRing, ph = (t / kRing) & 1;
      mbar_wait(&bars->kv_empty[st], ph ^ 1);
      if (lane == 0) {
        mbar_expect_tx(&bars->kv_full[st], kTileBytes);
        tma_load_2d(kv_smem + st * kTileBytes, &map_qkv, &bars->kv_full[st], ((t & 1) + 1) * p.C + h * kD, b * p.T + (t >> 1) * kBK);
      }
      __syncwarp();
    }
  } else if (warp == 5) {
    // ---- MMA issuer
    const uint32_t idesc_s = make_idesc(false, false, false, kBQ, kBK);          // S: A = Q (K-major), B = K (K-major)
    const uint32_t idesc_pv = make_idesc(false, false, true, kBQ, kD);           // PV: A = P (K-major), B = V (MN-major)
    const uint32_t qa = smem_u32(q_smem), pa = smem_u32(p_smem);
    auto issue_s = [&](int j) {
      const int t = 2 * j, st = t % kRing, ph = (t / kRing) & 1;
      mbar_wait(&bars->kv_full[st], ph);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t ka = smem_u32(kv_smem + st * kTileBytes);
#pragma unroll
        for (int k = 0; k < kD / 16; ++k)
          umma<false>(tS0 + (j & 1) * 128, make_smem_desc(qa + k * 32, 0, 1024), make_smem_desc(ka + k * 32, 0, 1024), idesc_s, k > 0);
        umma_commit(&bars->kv_empty[st]);      // K_j consumed
        umma_commit(&bars->s_full[j & 1]);
      }
      __syncwarp();
    };
    mbar_wait(&bars->q_full, 0);
    issue_s(0);
    for (int j = 0; j < nkv; ++j) {
      // p_ready(j): the softmax has consumed S_j, written P_j, and read P_{j-1} V_{j-1} out of the other buffer
      mbar_wait(&bars->p_ready, j & 1);
      if (j + 1 < nkv) issue_s(j + 1);         // into the other buffer, which is free now
      const int t = 2 * j + 1, st = t % kRing, ph = (t / kRing) & 1;
      mbar_wait(&bars->kv_full[st], ph);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t va = smem_u32(kv_smem + st * kTileBytes);
#pragma unroll
        for (int k = 0; k < kBK / 16; ++k) {
          // A: P chunk (k / 4) of 64 kv columns, 32 bytes per K step inside it; B: V rows 16 k .. 16 k + 15 (MN-major: 2048 B per step)
          const uint64_t da = make_smem_desc(pa + (k >> 2) * kTileBytes + (k & 3) * 32, 0, 1024);
          const uint64_t db = make_smem_desc(va + k * 2048, kBK * 128, 1024);
          umma<false>(tS0 + (j & 1) * 128, da, db, idesc_pv, k > 0);
        }
        umma_commit(&bars->kv_empty[st]);      // V_j consumed
        umma_commit(&bars->pv_full);
      }
      __syncwarp();
    }
  } else {
    // ---- softmax warps: thread = query row
    const int row = warp * 32 + lane;                        // row inside the block = TMEM lane
    const uint32_t lane_addr = (uint32_t)(warp * 32) << 16;
    float o[kD];
#pragma unroll
    for (int i = 0; i < kD; ++i) o[i] = 0.f;
    float m = -INFINITY, l = 0.f;
    for (int j = 0; j < nkv; ++j) {
      const uint32_t tS = tS0 + (j & 1) * 128 + lane_addr;
      const bool diag = j == qi;
      mbar_wait(&bars->s_full[j & 1], (j >> 1) & 1);
      tc_fence_after();
      // pass 1: row maximum
      float mx = -INFINITY;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        uint32_t r[32];
        tmem_ld32(tS + c * 32, r);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          const float v = __uint_as_float(r[i]);
          if (!diag || c * 32 + i <= row) mx = fmaxf(mx, v);
        }
      }
      const float m_new = fmaxf(m, mx * p.scale_log2);
      // previous block's P V arrives: add it, then rescale everything to the new maximum
      if (j > 0) {
        mbar_wait(&bars->pv_full, (j - 1) & 1);
        tc_fence_after();
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          uint32_t r[32];
          tmem_ld32(tS0 + ((j - 1) & 1) * 128 + lane_addr + c * 32, r);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) o[c * 32 + i] += __uint_as_float(r[i]);
        }
      }
      const float alpha = ex2(m - m_new);                    // m = -inf on the first block: alpha = 0, o and l are 0 anyway
      m = m_new;
      l *= alpha;
#pragma unroll
      for (int i = 0; i < kD; ++i) o[i] *= alpha;
      // pass 2: P = exp2(S * scale - m), bf16, into the swizzled K-major tile (previous P V has finished reading it: pv_full above)
      float sum = 0.f;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        uint32_t r[32];
        tmem_ld32(tS + c * 32, r);
        tmem_ld_wait();
        uint32_t pk[16];
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          float p0 = ex2(__uint_as_float(r[i]) * p.scale_log2 - m), p1 = ex2(__uint_as_float(r[i + 1]) * p.scale_log2 - m);
          if (diag) { if (c * 32 + i > row) p0 = 0.f; if (c * 32 + i + 1 > row) p1 = 0.f; }
          // sum what the MMA will see (bf16-rounded), so that the normalisation matches the numerator
          const __nv_bfloat162 q2 = __floats2bfloat162_rn(p0, p1);
          sum += __low2float(q2) + __high2float(q2);
          pk[i >> 1] = *reinterpret_cast<const uint32_t*>(&q2);
        }
        // 32 kv columns = 64 bytes = 4 units of 16 B: units (c & 1) * 4 .. + 3 of chunk (c >> 1)
        uint8_t* chunk = p_smem + (c >> 1) * kTileBytes;
#pragma unroll
        for (int u = 0; u < 4; ++u)
          *reinterpret_cast<uint4*>(chunk + swz(row, (c & 1) * 4 + u)) = make_uint4(pk[4 * u], pk[4 * u + 1], pk[4 * u + 2], pk[4 * u + 3]);
      }
      l += sum;
      tc_fence_before();
      fence_async_smem();                                    // generic-proxy stores -> visible to the tensor core's async proxy
      mbar_arrive(&bars->p_ready);
    }
    // last block's P V
    mbar_wait(&bars->pv_full, (nkv - 1) & 1);
    tc_fence_after();
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      uint32_t r[32];
      tmem_ld32(tS0 + ((nkv - 1) & 1) * 128 + lane_addr + c * 32, r);
      tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 32; ++i) o[c * 32 + i] += __uint_as_float(r[i]);
    }
    const float inv = 1.f / l;
    __nv_bfloat16* dst = p.out + (size_t)(row0 + row) * p.C + h * kD;
#pragma unroll
    for (int i = 0; i < kD; i += 8) {
      uint4 v = make_uint4(pack_bf16(o[i] * inv, o[i + 1] * inv), pack_bf16(o[i + 2] * inv, o[i + 3] * inv), pack_bf16(o[i + 4] * inv, o[i + 5] * inv),
                           pack_bf16(o[i + 6] * inv, o[i + 7] * inv));
      *reinterpret_cast<uint4*>(dst + i) = v;
    }
    p.lse[(size_t)bh * p.T + qi * kBQ + row] = m + log2f(l);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 5) tmem_dealloc<256>(tmem);
}

}  // namespace

extern "C" {

// Causal self-attention forward for head dimension 64.  qkv: bf16 [B*T, 3*H*64] (the fused projection: q | k | v, heads inside
// each third); out: bf16 [B*T, H*64]; lse: fp32 [B*H, T], log2-domain log-sum-exp of the scaled scores (consumed by
// dcp_attn_bwd).  T must be a multiple of 128.  softmax scale = 1/sqrt(64).
int dcp_attn_fwd(const void* qkv, void* out, float* lse, int B, int T, int H, cudaStream_t s) {
  if (T % kBQ != 0 || B < 1 || H < 1) { dcp_set_error("attention: T=%d must be a multiple of %d", T, kBQ); return -1; }
  const int C = H * kD;
  CUtensorMap map;
  if (encode_bf16_2d(&map, qkv, (long)B * T, 3L * C, 3L * C, kD, kBQ)) return -1;
  AttnFwdParams p{};
  p.T = T; p.H = H; p.C = C; p.scale_log2 = 0.125f * 1.4426950408889634f;
  p.out = static_cast<__nv_bfloat16*>(out); p.lse = lse;
  const size_t smem = (size_t)kTileBytes * (1 + kRing + 2) + sizeof(AttnBars) + 1024;
  static bool attr = false;
  if (!attr) { DCP_CUDA_CHECK(cudaFuncSetAttribute(k_attn_fwd, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); attr = true; }
  DCP_CUDA_CHECK(dcp_launch_pdl(k_attn_fwd, dim3(T / kBQ, B * H), dim3(kAttnThreads), smem, s, map, p));
  return 0;
}

}  // extern "C"
