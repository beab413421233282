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
//   warps 0-7 softmax:     TWO threads per query row (warps w and w + 4 own the same TMEM lanes): each takes 64 of the 128 key
//                          columns of S_j - one TMEM read into registers -, the two row maxima meet through shared memory, then
//                          P_j (bf16) goes straight into the 128-byte-swizzled K-major tile the PV MMA reads.  The output row
//                          lives in REGISTERS (32 floats per thread): each PV_j is added from TMEM and rescaled there, so
//                          TMEM never has to be rewritten.  Online softmax in the log2 domain.
// The row-wise log-sum-exp (log2 domain) is saved for the backward pass.
#include <cudaTypedefs.h>

#include "ops/tc_common.cuh"

namespace {
using namespace tc;

constexpr int kD = 64, kBQ = 128, kBK = 128;
constexpr int kAttnThreads = 320;                  // warps 0-7 compute (two threads per row), warp 8 TMA, warp 9 MMA
constexpr int kTileBytes = kBQ * kD * 2;            // 16 KB: one 128 x 64 bf16 tile (rows of 128 B, 128-byte swizzle)
constexpr int kRing = 3;                            // K / V tiles in flight

struct __align__(16) AttnBars {
  uint64_t q_full, kv_full[kRing], kv_empty[kRing], s_full[2], p_ready, pv_full;
  uint32_t tmem_base;
  float xmax[2][2][kBQ];       // [block parity][column half][row]: row maxima of the two column halves
  float xsum[kBQ];             // second half's row sum, for the final normalisation
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
  if (warp == 8 && lane == 0) tma_prefetch_desc(&map_qkv);
  if (warp == 9) {
    if (lane == 0) {
      mbar_init(&bars->q_full, 1);
      for (int i = 0; i < kRing; ++i) { mbar_init(&bars->kv_full[i], 1); mbar_init(&bars->kv_empty[i], 1); }
      for (int i = 0; i < 2; ++i) mbar_init(&bars->s_full[i], 1);
      mbar_init(&bars->p_ready, 256); mbar_init(&bars->pv_full, 1);
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

  if (warp == 8) {
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
  } else if (warp == 9) {
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
    // ---- softmax warps: two threads per query row; `half` selects the 64 key columns (and the 32 output columns) of this thread
    const int half = warp >> 2, lg = warp & 3;
    const int row = lg * 32 + lane;                          // row inside the block = TMEM lane
    const uint32_t lane_addr = (uint32_t)(lg * 32) << 16;
    float o[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) o[i] = 0.f;
    float m = -INFINITY, l = 0.f;
    for (int j = 0; j < nkv; ++j) {
      const uint32_t tS = tS0 + (j & 1) * 128 + lane_addr + half * 64;
      const bool diag = j == qi;
      mbar_wait(&bars->s_full[j & 1], (j >> 1) & 1);
      tc_fence_after();
      // pass 1: maximum of this thread's 64 columns (S_j is read again below: TMEM reads are cheap, 64 live registers are not
      // when two CTAs share an SM)
      float mx = -INFINITY;
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        uint32_t r[32];
        tmem_ld32(tS + c * 32, r);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i)
          if (!diag || half * 64 + c * 32 + i <= row) mx = fmaxf(mx, __uint_as_float(r[i]));
      }
      bars->xmax[j & 1][half][row] = mx;
      asm volatile("bar.sync 1, 256;" ::: "memory");
      const float m_new = fmaxf(m, fmaxf(mx, bars->xmax[j & 1][half ^ 1][row]) * p.scale_log2);
      // previous block's P V arrives: add it, then rescale everything to the new maximum
      if (j > 0) {
        mbar_wait(&bars->pv_full, (j - 1) & 1);
        tc_fence_after();
        uint32_t r[32];
        tmem_ld32(tS0 + ((j - 1) & 1) * 128 + lane_addr + half * 32, r);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) o[i] += __uint_as_float(r[i]);
      }
      const float alpha = ex2(m - m_new);                    // m = -inf on the first block: alpha = 0, o and l are 0 anyway
      m = m_new;
      l *= alpha;
#pragma unroll
      for (int i = 0; i < 32; ++i) o[i] *= alpha;
      // P = exp2(S * scale - m), bf16, into chunk `half` of the swizzled K-major tile (the previous P V has finished reading it)
      float sum = 0.f;
      uint8_t* chunk = p_smem + half * kTileBytes;
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        uint32_t r[32];
        tmem_ld32(tS + c * 32, r);
        tmem_ld_wait();
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          uint32_t pk[4];
#pragma unroll
          for (int i = 0; i < 8; i += 2) {
            const int cc = c * 32 + u * 8 + i;
            float p0 = ex2(__uint_as_float(r[u * 8 + i]) * p.scale_log2 - m), p1 = ex2(__uint_as_float(r[u * 8 + i + 1]) * p.scale_log2 - m);
            if (diag) { if (half * 64 + cc > row) p0 = 0.f; if (half * 64 + cc + 1 > row) p1 = 0.f; }
            // sum what the MMA will see (bf16-rounded), so that the normalisation matches the numerator
            const __nv_bfloat162 q2 = __floats2bfloat162_rn(p0, p1);
            sum += __low2float(q2) + __high2float(q2);
            pk[i >> 1] = *reinterpret_cast<const uint32_t*>(&q2);
          }
          *reinterpret_cast<uint4*>(chunk + swz(row, c * 4 + u)) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
        }
      }
      l += sum;
      tc_fence_before();
      fence_async_smem();                                    // generic-proxy stores -> visible to the tensor core's async proxy
      mbar_arrive(&bars->p_ready);
    }
    // last block's P V
    mbar_wait(&bars->pv_full, (nkv - 1) & 1);
    tc_fence_after();
    {
      uint32_t r[32];
      tmem_ld32(tS0 + ((nkv - 1) & 1) * 128 + lane_addr + half * 32, r);
      tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 32; ++i) o[i] += __uint_as_float(r[i]);
    }
    if (half == 1) bars->xsum[row] = l;
    asm volatile("bar.sync 1, 256;" ::: "memory");
    if (half == 0) l += bars->xsum[row];
    else l += 0.f;
    if (half == 0) bars->xsum[row] = l;
    asm volatile("bar.sync 1, 256;" ::: "memory");
    l = bars->xsum[row];
    const float inv = 1.f / l;
    __nv_bfloat16* dst = p.out + (size_t)(row0 + row) * p.C + h * kD + half * 32;
#pragma unroll
    for (int i = 0; i < 32; i += 8) {
      uint4 v = make_uint4(pack_bf16(o[i] * inv, o[i + 1] * inv), pack_bf16(o[i + 2] * inv, o[i + 3] * inv), pack_bf16(o[i + 4] * inv, o[i + 5] * inv),
                           pack_bf16(o[i + 6] * inv, o[i + 7] * inv));
      *reinterpret_cast<uint4*>(dst + i) = v;
    }
    if (half == 0) p.lse[(size_t)bh * p.T + qi * kBQ + row] = m + log2f(l);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 9) tmem_dealloc<256>(tmem);
}

// =====================================================================================================
// Backward.  One CTA per (128-row key/value block j, batch*head); query blocks i = j .. T/128 - 1 stream through it.
//   S  = Q_i K_j^T            dP = dO_i V_j^T                         (two 128x128x64 MMAs into TMEM)
//   P  = exp2(S * scale - lse_i)          dS = P o (dP - delta_i) * softmax_scale      (threads: row = query, written as bf16 into
//                                                                                         two swizzled [query][key] tiles)
//   dV += P^T dO_i            dK += dS^T Q_i            dQ_i = dS K_j  (three 128x64x128 MMAs; P / dS are read both MN-major - the
//                                                                        transposed products - and K-major from the same tiles)
// dV / dK accumulate in TMEM over the whole loop and are written once (bf16, straight into the dqkv matrix next to dq, i.e. in
// the layout the qkv projection's backward GEMMs read); dQ_i is added to an fp32 accumulator with vector reductions
// (red.global.add.v4.f32: several key blocks contribute to one query block) and converted by k_attn_dq_convert.
//   warp 4 TMA: K_j, V_j once; (Q_i, dO_i) through two stages        warp 5 MMA issuer        warps 0-3: thread = query row
// =====================================================================================================
struct __align__(16) AttnBwdBars {
  uint64_t kv_full, qdo_full[2], qdo_empty[2], sdp_full, pds_ready, pds_free, dq_full, dq_free;
  uint32_t tmem_base;
};

struct AttnBwdParams {
  int T, H, C;
  float scale_log2, scale;
  const float* lse;            // [B*H, T] log2 domain
  const float* delta;          // [B*H, T] rowsum(dO o O)
  float* dq_acc;               // [B*T, C] fp32, zeroed by the caller
  __nv_bfloat16* dqkv;         // [B*T, 3C]: dk / dv written here (columns C + h*64, 2C + h*64)
};

__device__ __forceinline__ void red_add_v4(float* p, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

__global__ void __launch_bounds__(kAttnThreads, 1)
k_attn_bwd(const __grid_constant__ CUtensorMap map_qkv, const __grid_constant__ CUtensorMap map_do, const AttnBwdParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = align_smem_1024(smem_raw);
  uint8_t* k_smem = smem;                                   // 16 KB
  uint8_t* v_smem = smem + kTileBytes;                      // 16 KB
  uint8_t* qdo_smem = smem + 2 * kTileBytes;                // 2 stages x (Q 16 KB + dO 16 KB)
  uint8_t* p_smem = qdo_smem + 4 * kTileBytes;              // 32 KB: P  [query][key], two 64-key chunks
  uint8_t* ds_smem = p_smem + 2 * kTileBytes;               // 32 KB: dS [query][key]
  AttnBwdBars* bars = reinterpret_cast<AttnBwdBars*>(ds_smem + 2 * kTileBytes);
  const int warp = warp_idx_uniform(), lane = threadIdx.x & 31;
  const int kj = blockIdx.x;                                // key/value block: 0 (most query blocks) first
  const int bh = blockIdx.y, b = bh / p.H, h = bh % p.H;
  const int nq = p.T / kBQ, nit = nq - kj;

  griddep_launch();
  if (warp == 8 && lane == 0) { tma_prefetch_desc(&map_qkv); tma_prefetch_desc(&map_do); }
  if (warp == 9) {
    if (lane == 0) {
      mbar_init(&bars->kv_full, 1);
      for (int i = 0; i < 2; ++i) { mbar_init(&bars->qdo_full[i], 1); mbar_init(&bars->qdo_empty[i], 1); }
      mbar_init(&bars->sdp_full, 1); mbar_init(&bars->pds_ready, 256); mbar_init(&bars->pds_free, 1);
      mbar_init(&bars->dq_full, 1); mbar_init(&bars->dq_free, 256);
      mbar_fence_init();
    }
    __syncwarp();
    tmem_alloc<512>(&bars->tmem_base);
  }
  griddep_wait();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *((volatile uint32_t*)&bars->tmem_base);
  const uint32_t tS = tmem, tDP = tmem + 128, tDV = tmem + 256, tDK = tmem + 320, tDQ = tmem + 384;

  if (warp == 8) {
    if (lane == 0) {
      mbar_expect_tx(&bars->kv_full, 2 * kTileBytes);
      tma_load_2d(k_smem, &map_qkv, &bars->kv_full, p.C + h * kD, b * p.T + kj * kBK);
      tma_load_2d(v_smem, &map_qkv, &bars->kv_full, 2 * p.C + h * kD, b * p.T + kj * kBK);
    }
    __syncwarp();
    for (int it = 0; it < nit; ++it) {
      const int st = it & 1, ph = (it >> 1) & 1;
      mbar_wait(&bars->qdo_empty[st], ph ^ 1);
      if (lane == 0) {
        const int r0 = b * p.T + (kj + it) * kBQ;
        mbar_expect_tx(&bars->qdo_full[st], 2 * kTileBytes);
        tma_load_2d(qdo_smem + st * 2 * kTileBytes, &map_qkv, &bars->qdo_full[st], h * kD, r0);
        tma_load_2d(qdo_smem + st * 2 * kTileBytes + kTileBytes, &map_do, &bars->qdo_full[st], h * kD, r0);
      }
      __syncwarp();
    }
  } else if (warp == 9) {
    const uint32_t idesc_s = make_idesc(false, false, false, kBQ, kBK);      // S / dP: both operands K-major, 128 x 128
    const uint32_t idesc_t = make_idesc(false, true, true, kBK, kD);         // dV / dK: A = P^T / dS^T (MN-major), B = dO / Q (MN-major)
    const uint32_t idesc_q = make_idesc(false, false, true, kBQ, kD);        // dQ: A = dS (K-major), B = K (MN-major)
    const uint32_t ka = smem_u32(k_smem), va = smem_u32(v_smem), pa = smem_u32(p_smem), dsa = smem_u32(ds_smem);
    auto issue_sdp = [&](int it) {
      const int st = it & 1;
      mbar_wait(&bars->qdo_full[st], (it >> 1) & 1);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t qa = smem_u32(qdo_smem + st * 2 * kTileBytes), doa = qa + kTileBytes;
#pragma unroll
        for (int k = 0; k < kD / 16; ++k)
          umma<false>(tS, make_smem_desc(qa + k * 32, 0, 1024), make_smem_desc(ka + k * 32, 0, 1024), idesc_s, k > 0);
#pragma unroll
        for (int k = 0; k < kD / 16; ++k)
          umma<false>(tDP, make_smem_desc(doa + k * 32, 0, 1024), make_smem_desc(va + k * 32, 0, 1024), idesc_s, k > 0);
        umma_commit(&bars->sdp_full);
      }
      __syncwarp();
    };
    mbar_wait(&bars->kv_full, 0);
    issue_sdp(0);
    for (int it = 0; it < nit; ++it) {
      const int st = it & 1;
      mbar_wait(&bars->pds_ready, it & 1);                 // P / dS of this iteration are in shared memory; S / dP have been consumed
      if (it + 1 < nit) issue_sdp(it + 1);                 // the next scores are computed while the three products below run
      if (it > 0) mbar_wait(&bars->dq_free, (it - 1) & 1); // the threads have taken dQ of the previous iteration
      tc_fence_after();
      if (elect_one()) {
        const uint32_t qa = smem_u32(qdo_smem + st * 2 * kTileBytes), doa = qa + kTileBytes;
#pragma unroll
        for (int k = 0; k < kBQ / 16; ++k) {
          // K index = query row: 16 rows = 2048 B per step in every operand here; LBO = next 64-key chunk of P / dS
          const uint64_t pT = make_smem_desc(pa + k * 2048, kTileBytes, 1024), dsT = make_smem_desc(dsa + k * 2048, kTileBytes, 1024);
          umma<false>(tDV, pT, make_smem_desc(doa + k * 2048, kTileBytes, 1024), idesc_t, (it > 0) || (k > 0));
          umma<false>(tDK, dsT, make_smem_desc(qa + k * 2048, kTileBytes, 1024), idesc_t, (it > 0) || (k > 0));
        }
#pragma unroll
        for (int k = 0; k < kBK / 16; ++k)                 // K index = key: dS chunk (k / 4), 32 B per step; K_j rows 16 k ..
          umma<false>(tDQ, make_smem_desc(dsa + (k >> 2) * kTileBytes + (k & 3) * 32, 0, 1024), make_smem_desc(ka + k * 2048, kTileBytes, 1024), idesc_q, k > 0);
        umma_commit(&bars->qdo_empty[st]);
        umma_commit(&bars->pds_free);
        umma_commit(&bars->dq_full);
      }
      __syncwarp();
    }
  } else {
    // two threads per query row: `half` selects the 64 key columns of S / dP (= chunk `half` of the P / dS tiles), the 32 dQ
    // columns, and at the end which of dK / dV this thread writes
    const int half = warp >> 2, lg = warp & 3;
    const int row = lg * 32 + lane;
    const uint32_t lane_addr = (uint32_t)(lg * 32) << 16;
    for (int it = 0; it < nit; ++it) {
      const int qi = kj + it;
      const bool diag = it == 0;
      const size_t grow = (size_t)bh * p.T + qi * kBQ + row;
      const float lse = p.lse[grow], delta = p.delta[grow];
      mbar_wait(&bars->sdp_full, it & 1);
      if (it > 0) mbar_wait(&bars->pds_free, (it - 1) & 1);  // the previous iteration's products no longer read the P / dS tiles
      tc_fence_after();
      uint8_t* pc = p_smem + half * kTileBytes;
      uint8_t* dc = ds_smem + half * kTileBytes;
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        uint32_t rs[32], rp[32];
        tmem_ld32(tS + lane_addr + half * 64 + c * 32, rs);
        tmem_ld32(tDP + lane_addr + half * 64 + c * 32, rp);
        tmem_ld_wait();
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          uint32_t pk[4], dk[4];
#pragma unroll
          for (int i = 0; i < 8; i += 2) {
            const int e = u * 8 + i, col = half * 64 + c * 32 + e;
            float p0 = ex2(__uint_as_float(rs[e]) * p.scale_log2 - lse), p1 = ex2(__uint_as_float(rs[e + 1]) * p.scale_log2 - lse);
            if (diag) { if (col > row) p0 = 0.f; if (col + 1 > row) p1 = 0.f; }
            const float d0 = p0 * (__uint_as_float(rp[e]) - delta) * p.scale, d1 = p1 * (__uint_as_float(rp[e + 1]) - delta) * p.scale;
            pk[i >> 1] = pack_bf16(p0, p1);
            dk[i >> 1] = pack_bf16(d0, d1);
          }
          *reinterpret_cast<uint4*>(pc + swz(row, c * 4 + u)) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
          *reinterpret_cast<uint4*>(dc + swz(row, c * 4 + u)) = make_uint4(dk[0], dk[1], dk[2], dk[3]);
        }
      }
      tc_fence_before();
      fence_async_smem();
      mbar_arrive(&bars->pds_ready);
      // dQ_i of this key block -> fp32 accumulator (32 of the 64 columns per thread)
      mbar_wait(&bars->dq_full, it & 1);
      tc_fence_after();
      float* dq = p.dq_acc + ((size_t)b * p.T + qi * kBQ + row) * p.C + h * kD + half * 32;
      {
        uint32_t r[32];
        tmem_ld32(tDQ + lane_addr + half * 32, r);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; i += 4)
          red_add_v4(dq + i, __uint_as_float(r[i]), __uint_as_float(r[i + 1]), __uint_as_float(r[i + 2]), __uint_as_float(r[i + 3]));
      }
      tc_fence_before();
      mbar_arrive(&bars->dq_free);
    }
    // dK (half 0) / dV (half 1) of this key block: thread = key row
    mbar_wait(&bars->pds_free, (nit - 1) & 1);
    tc_fence_after();
    __nv_bfloat16* dst = p.dqkv + ((size_t)b * p.T + kj * kBK + row) * (3 * p.C) + h * kD + (half + 1) * p.C;
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      uint32_t r[32];
      tmem_ld32((half ? tDV : tDK) + lane_addr + c * 32, r);
      tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 32; i += 8)
        *reinterpret_cast<uint4*>(dst + c * 32 + i) = make_uint4(pack_bf16(__uint_as_float(r[i]), __uint_as_float(r[i + 1])), pack_bf16(__uint_as_float(r[i + 2]), __uint_as_float(r[i + 3])),
                                                                  pack_bf16(__uint_as_float(r[i + 4]), __uint_as_float(r[i + 5])), pack_bf16(__uint_as_float(r[i + 6]), __uint_as_float(r[i + 7])));
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 9) tmem_dealloc<512>(tmem);
}

// delta[bh, t] = sum_d dO[t, h, d] * O[t, h, d]: one thread per (row, head), 64 products
__global__ void __launch_bounds__(256) k_attn_delta(const __nv_bfloat16* __restrict__ o, const __nv_bfloat16* __restrict__ dout, float* __restrict__ delta,
                                                    int BT, int T, int H) {
  griddep_launch();
  griddep_wait();
  const long idx = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long)BT * H) return;
  const int h = (int)(idx % H); const long r = idx / H;
  const uint4* po = reinterpret_cast<const uint4*>(o + r * (long)H * kD + h * kD);
  const uint4* pd = reinterpret_cast<const uint4*>(dout + r * (long)H * kD + h * kD);
  float acc = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const uint4 a = po[i], d = pd[i];
    const __nv_bfloat162* a2 = reinterpret_cast<const __nv_bfloat162*>(&a);
    const __nv_bfloat162* d2 = reinterpret_cast<const __nv_bfloat162*>(&d);
#pragma unroll
    for (int k = 0; k < 4; ++k) acc += __low2float(a2[k]) * __low2float(d2[k]) + __high2float(a2[k]) * __high2float(d2[k]);
  }
  const long b = r / T, t = r % T;
  delta[(b * H + h) * T + t] = acc;
}

// dq (fp32 accumulator [B*T, C]) -> bf16 into columns [0, C) of dqkv [B*T, 3C]
__global__ void __launch_bounds__(256) k_attn_dq_convert(const float* __restrict__ acc, __nv_bfloat16* __restrict__ dqkv, long rows, int C) {
  griddep_launch();
  griddep_wait();
  const long n8 = rows * C / 8;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (long)gridDim.x * blockDim.x) {
    const long r = (i * 8) / C; const int c = (int)((i * 8) % C);
    const float4 a = *reinterpret_cast<const float4*>(acc + i * 8), b = *reinterpret_cast<const float4*>(acc + i * 8 + 4);
    *reinterpret_cast<uint4*>(dqkv + r * 3L * C + c) = make_uint4(pack_bf16(a.x, a.y), pack_bf16(a.z, a.w), pack_bf16(b.x, b.y), pack_bf16(b.z, b.w));
  }
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

// Backward of dcp_attn_fwd.  dout / out: bf16 [B*T, H*64]; lse from the forward; dqkv: bf16 [B*T, 3*H*64] receives dq | dk | dv
// in the layout of qkv (what the projection's backward GEMMs consume - no split / merge copies).  Scratch from the caller:
// delta (B*H*T floats) and dq_acc (B*T*H*64 floats; zeroed here, stream-ordered).
int dcp_attn_bwd(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv, float* delta, float* dq_acc, int B, int T, int H,
                 cudaStream_t s) {
  if (T % kBQ != 0 || B < 1 || H < 1) { dcp_set_error("attention: T=%d must be a multiple of %d", T, kBQ); return -1; }
  const int C = H * kD;
  const long BT = (long)B * T;
  CUtensorMap mq, md;
  if (encode_bf16_2d(&mq, qkv, BT, 3L * C, 3L * C, kD, kBQ)) return -1;
  if (encode_bf16_2d(&md, dout, BT, C, C, kD, kBQ)) return -1;
  DCP_CUDA_CHECK(cudaMemsetAsync(dq_acc, 0, sizeof(float) * BT * C, s));
  DCP_CUDA_CHECK(dcp_launch_pdl(k_attn_delta, dim3((unsigned)((BT * H + 255) / 256)), dim3(256), 0, s, static_cast<const __nv_bfloat16*>(out),
                                static_cast<const __nv_bfloat16*>(dout), delta, (int)BT, T, H));
  AttnBwdParams p{};
  p.T = T; p.H = H; p.C = C; p.scale = 0.125f; p.scale_log2 = 0.125f * 1.4426950408889634f;
  p.lse = lse; p.delta = delta; p.dq_acc = dq_acc; p.dqkv = static_cast<__nv_bfloat16*>(dqkv);
  const size_t smem = (size_t)kTileBytes * 10 + sizeof(AttnBwdBars) + 1024;
  static bool attr = false;
  if (!attr) { DCP_CUDA_CHECK(cudaFuncSetAttribute(k_attn_bwd, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); attr = true; }
  DCP_CUDA_CHECK(dcp_launch_pdl(k_attn_bwd, dim3(T / kBK, B * H), dim3(kAttnThreads), smem, s, mq, md, p));
  DCP_CUDA_CHECK(dcp_launch_pdl(k_attn_dq_convert, dim3(148 * 4), dim3(256), 0, s, static_cast<const float*>(dq_acc), static_cast<__nv_bfloat16*>(dqkv), BT, C));
  return 0;
}

}  // extern "C"
