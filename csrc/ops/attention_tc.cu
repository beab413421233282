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
#include <cstdlib>
#include <type_traits>
#include <cudaTypedefs.h>

#include "ops/tc_common.cuh"

namespace {
using namespace tc;

constexpr int kD = 64, kBQ = 128, kBK = 128;
constexpr int kAttnThreads = 320;                  // warps 0-7 compute (two threads per row), warp 8 TMA, warp 9 MMA
constexpr int kTileBytes = kBQ * kD * 2;            // 16 KB: one 128 x 64 bf16 tile (rows of 128 B, 128-byte swizzle)
constexpr int kRing = 3;                            // K / V tiles in flight

struct __align__(16) AttnBars {
  uint64_t q_full, q_empty, kv_full[kRing], kv_empty[kRing], s_full[2], o_taken, p_ready, pv_full;
  uint32_t tmem_base;
  float xmax[2][2][kBQ];       // [block parity][column half][row]: row maxima of the two column halves
  float xsum[kBQ];             // second half's row sum, for the final normalisation
};

struct AttnFwdParams {
  int T, H, C, BH;             // sequence length, heads, model width (H * 64), batch * heads
  float scale_log2;            // softmax scale * log2(e)
  __nv_bfloat16* out;          // [B*T, C]
  float* lse;                  // [B*H, T] log2-domain log-sum-exp
  long long* trace;            // debugging: clock64 stamps of CTA (trace_x, 0) (dcp_attn_set_trace), else nullptr
  int trace_x;
};
#define ATTN_FSTAMP(j, k) do { if (p.trace != nullptr && (j) < 32 && blockIdx.x == p.trace_x && lane == 0 && (warp == 0 || warp == 9)) p.trace[(j) * 8 + (k)] = clock64(); } while (0)

// 16-byte unit `u` (0..7) of row r inside a 128-row x 128-byte swizzled chunk
__device__ __forceinline__ uint32_t swz(uint32_t r, uint32_t u) { return r * 128u + ((u ^ (r & 7u)) << 4); }

__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
// bulk-group side of the copy engine: tile store through a tensor map, fp32 reduce-add of a contiguous block
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* m, const void* src_smem, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(m), "r"(smem_u32(src_smem)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void bulk_reduce_add_f32(float* dst, const void* src_smem, uint32_t bytes) {
  asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], %2;" ::"l"(dst), "r"(smem_u32(src_smem)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait0() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ float ex2(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
// packed fp32 pairs (FFMA2 / FMUL2 / FADD2) and the three-input maximum (FMNMX3): half the issue slots of the scalar forms
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
  float2 d;
  asm("fma.rn.ftz.f32x2 %0, %1, %2, %3;" : "=l"(reinterpret_cast<uint64_t&>(d)) : "l"(reinterpret_cast<const uint64_t&>(a)), "l"(reinterpret_cast<const uint64_t&>(b)), "l"(reinterpret_cast<const uint64_t&>(c)));
  return d;
}
__device__ __forceinline__ float2 fmul2(float2 a, float2 b) {
  float2 d;
  asm("mul.rn.ftz.f32x2 %0, %1, %2;" : "=l"(reinterpret_cast<uint64_t&>(d)) : "l"(reinterpret_cast<const uint64_t&>(a)), "l"(reinterpret_cast<const uint64_t&>(b)));
  return d;
}
__device__ __forceinline__ float2 fadd2(float2 a, float2 b) {
  float2 d;
  asm("add.rn.ftz.f32x2 %0, %1, %2;" : "=l"(reinterpret_cast<uint64_t&>(d)) : "l"(reinterpret_cast<const uint64_t&>(a)), "l"(reinterpret_cast<const uint64_t&>(b)));
  return d;
}
__device__ __forceinline__ float fmax3(float a, float b, float c) { float d; asm("max.ftz.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c)); return d; }
__device__ __forceinline__ float2 u2f2(uint32_t a, uint32_t b) { return make_float2(__uint_as_float(a), __uint_as_float(b)); }
__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}

// the r-th item of CTA c out of G: rounds alternate direction (c, 2G-1-c, 2G+c, ...); items are ordered heaviest first, so
// every CTA collects a similar number of tiles
__device__ __forceinline__ int attn_item(int r, int c, int G) { return r * G + ((r & 1) ? G - 1 - c : c); }

// PERSISTENT: two CTAs per SM, each walking its list of (query block, batch*head) items; the key blocks of consecutive items
// form ONE stream (counter g) for all three roles, so the producer fetches the next item's Q / K_0 while the last block's
// softmax runs and the MMA warp issues its first S ahead of the current item's last P V.
__global__ void __launch_bounds__(kAttnThreads, 2)
k_attn_fwd(const __grid_constant__ CUtensorMap map_qkv, const __grid_constant__ CUtensorMap map_out, const AttnFwdParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = align_smem_1024(smem_raw);
  uint8_t* q_smem = smem;                                   // 16 KB
  uint8_t* kv_smem = smem + kTileBytes;                     // kRing x 16 KB: K_0, V_0, K_1, V_1, ...
  uint8_t* p_smem = kv_smem + kRing * kTileBytes;           // 32 KB: P as two 64-column K-major chunks
  AttnBars* bars = reinterpret_cast<AttnBars*>(p_smem + 2 * kTileBytes);
  const int warp = warp_idx_uniform(), lane = threadIdx.x & 31;
  const int nq = p.T / kBQ, nitems = nq * p.BH, G = gridDim.x, cta = blockIdx.x;
  // item w: query block nq - 1 - w / BH (late = heavy query blocks first), batch*head w % BH

  ATTN_FSTAMP(30, 0);
  griddep_launch();
  if (warp == 8 && lane == 0) { tma_prefetch_desc(&map_qkv); tma_prefetch_desc(&map_out); }
  if (warp == 9) {
    if (lane == 0) {
      mbar_init(&bars->q_full, 1); mbar_init(&bars->q_empty, 1);
      for (int i = 0; i < kRing; ++i) { mbar_init(&bars->kv_full[i], 1); mbar_init(&bars->kv_empty[i], 1); }
      for (int i = 0; i < 2; ++i) mbar_init(&bars->s_full[i], 1);
      mbar_init(&bars->o_taken, 256); mbar_init(&bars->p_ready, 256); mbar_init(&bars->pv_full, 1);
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
  ATTN_FSTAMP(30, 1);
  const uint32_t tS0 = tmem;                                // S_g (128 columns) and later P_g V_g (64 columns) in buffer g & 1

  if (warp == 8) {
    // ---- TMA producer
    int t = 0;                                              // K / V tile stream: K of block g is tile 2 g, V is tile 2 g + 1
    for (int n = 0, w = attn_item(0, cta, G); w < nitems; ++n, w = attn_item(n, cta, G)) {
      const int qi = nq - 1 - w / p.BH, bh = w % p.BH, b = bh / p.H, h = bh % p.H;
      mbar_wait(&bars->q_empty, (n & 1) ^ 1);               // the previous item's last S has been computed
      if (lane == 0) {
        mbar_expect_tx(&bars->q_full, kTileBytes);
        tma_load_2d(q_smem, &map_qkv, &bars->q_full, h * kD, b * p.T + qi * kBQ);
      }
      __syncwarp();
      for (int tt = 0; tt < 2 * (qi + 1); ++tt, ++t) {
        const int st = t % kRing, ph = (t / kRing) & 1;
        mbar_wait(&bars->kv_empty[st], ph ^ 1);
        if (lane == 0) {
          mbar_expect_tx(&bars->kv_full[st], kTileBytes);
          tma_load_2d(kv_smem + st * kTileBytes, &map_qkv, &bars->kv_full[st], ((tt & 1) + 1) * p.C + h * kD, b * p.T + (tt >> 1) * kBK);
        }
        __syncwarp();
      }
    }
  } else if (warp == 9) {
    // ---- MMA issuer
    const uint32_t idesc_s = make_idesc(false, false, false, kBQ, kBK);          // S: A = Q (K-major), B = K (K-major)
    const uint32_t idesc_pv = make_idesc(false, false, true, kBQ, kD);           // PV: A = P (K-major), B = V (MN-major)
    const uint32_t qa = smem_u32(q_smem), pa = smem_u32(p_smem);
    // S of stream block g; `last`: the item's last key block - Q may be replaced once this product is done
    auto issue_s = [&](int g, bool last) {
      const int t = 2 * g, st = t % kRing, ph = (t / kRing) & 1;
      mbar_wait(&bars->kv_full[st], ph);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t ka = smem_u32(kv_smem + st * kTileBytes);
#pragma unroll
        for (int k = 0; k < kD / 16; ++k)
          umma<false>(tS0 + (g & 1) * 128, make_smem_desc(qa + k * 32, 0, 1024), make_smem_desc(ka + k * 32, 0, 1024), idesc_s, k > 0);
        umma_commit(&bars->kv_empty[st]);      // K consumed
        umma_commit(&bars->s_full[g & 1]);
        if (last) umma_commit(&bars->q_empty);
      }
      __syncwarp();
    };
    int g = 0;
    if (attn_item(0, cta, G) < nitems) {
      mbar_wait(&bars->q_full, 0);
      issue_s(0, nq - 1 - attn_item(0, cta, G) / p.BH == 0);
    }
    for (int n = 0, w = attn_item(0, cta, G); w < nitems; ++n, w = attn_item(n, cta, G)) {
      const int nkv = nq - w / p.BH;
      const int w_next = attn_item(n + 1, cta, G);
      for (int j = 0; j < nkv; ++j, ++g) {
        // o_taken(g): the softmax threads have read P_{g-1} V_{g-1} out of the other buffer (early in block g, before its
        // exponentials): the next S (of this item or of the next one) goes there now and is ready when they finish block g
        mbar_wait(&bars->o_taken, g & 1);
        tc_fence_after();
        ATTN_FSTAMP(g, 4);
        if (j + 1 < nkv) issue_s(g + 1, j + 2 == nkv);
        else if (w_next < nitems) { mbar_wait(&bars->q_full, (n + 1) & 1); issue_s(g + 1, nq - 1 - w_next / p.BH == 0); }
        ATTN_FSTAMP(g, 5);
        // p_ready(g): S_g has been consumed and P_g written
        mbar_wait(&bars->p_ready, g & 1);
        const int t = 2 * g + 1, st = t % kRing, ph = (t / kRing) & 1;
        mbar_wait(&bars->kv_full[st], ph);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t va = smem_u32(kv_smem + st * kTileBytes);
#pragma unroll
          for (int k = 0; k < kBK / 16; ++k) {
            // A: P chunk (k / 4) of 64 kv columns, 32 bytes per K step inside it; B: V rows 16 k .. 16 k + 15 (MN-major: 2048 B per step)
            const uint64_t da = make_smem_desc(pa + (k >> 2) * kTileBytes + (k & 3) * 32, 0, 1024);
            const uint64_t db = make_smem_desc(va + k * 2048, kBK * 128, 1024);
            umma<false>(tS0 + (g & 1) * 128, da, db, idesc_pv, k > 0);
          }
          umma_commit(&bars->kv_empty[st]);      // V consumed
          umma_commit(&bars->pv_full);
        }
        __syncwarp();
        ATTN_FSTAMP(g, 6);
      }
    }
  } else {
    // ---- softmax warps: two threads per query row; `half` selects the 64 key columns (and the 32 output columns) of this thread
    const int half = warp >> 2, lg = warp & 3;
    const int row = lg * 32 + lane;                          // row inside the block = TMEM lane
    const uint32_t lane_addr = (uint32_t)(lg * 32) << 16;
    const float2 sc2 = make_float2(p.scale_log2, p.scale_log2);
    uint8_t* chunk = p_smem + half * kTileBytes;
    float2 o[16];
    float m, l;
    int g = 0;
    // one key block (stream index g, block j of its item); kDiag: the block on the diagonal, the only one that needs the causal mask
    auto tile = [&](auto dtag, int j, bool first_of_later_item) {
      constexpr bool kDiag = decltype(dtag)::value;
      const uint32_t tS = tS0 + (g & 1) * 128 + lane_addr + half * 64;
      mbar_wait(&bars->s_full[g & 1], (g >> 1) & 1);
      tc_fence_after();
      ATTN_FSTAMP(g, 0);
      // pass 1: maximum of this thread's 64 columns (S is read again below: TMEM reads are cheap, 64 live registers are not
      // when two CTAs share an SM)
      // On the diagonal block the 32-column chunk hc = 2 half + c of warp group lg (rows 32 lg ..) is entirely below the
      // diagonal for hc < lg, entirely above it for hc > lg (skipped: nothing to read, P = 0) and cut by it for hc == lg
      // (column i of the chunk is visible to lane >= i) - all warp-uniform, so the mask costs one chunk in four.
      float mx = -INFINITY;
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        const int hc = half * 2 + c;
        if (kDiag && hc > lg) continue;
        uint32_t r[32];
        tmem_ld32(tS + c * 32, r);
        tmem_ld_wait();
        if (kDiag && hc == lg) {
#pragma unroll
          for (int i = 0; i < 32; i += 2)
            mx = fmax3(mx, i > lane ? -INFINITY : __uint_as_float(r[i]), i + 1 > lane ? -INFINITY : __uint_as_float(r[i + 1]));
        } else {
#pragma unroll
          for (int i = 0; i < 32; i += 2) mx = fmax3(mx, __uint_as_float(r[i]), __uint_as_float(r[i + 1]));
        }
      }
      bars->xmax[g & 1][half][row] = mx;
      ATTN_FSTAMP(g, 1);
      if (first_of_later_item && threadIdx.x == 0) bulk_wait_read0();   // the previous item's output tile has left the P buffer
      asm volatile("bar.sync 1, 256;" ::: "memory");
      const float m_new = fmaxf(m, fmaxf(mx, bars->xmax[g & 1][half ^ 1][row]) * p.scale_log2);
      const float alpha = ex2(m - m_new);                    // m = -inf on the first block: alpha = 0, o and l are 0 anyway
      const float2 al2 = make_float2(alpha, alpha);
      // previous block's P V arrives: add it, then rescale everything to the new maximum
      if (j > 0) {
        mbar_wait(&bars->pv_full, (g - 1) & 1);
        tc_fence_after();
        uint32_t r[32];
        tmem_ld32(tS0 + ((g - 1) & 1) * 128 + lane_addr + half * 32, r);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 16; ++i) o[i] = fmul2(fadd2(o[i], u2f2(r[2 * i], r[2 * i + 1])), al2);
      }
      tc_fence_before();
      mbar_arrive(&bars->o_taken);
      m = m_new;
      l *= alpha;
      ATTN_FSTAMP(g, 2);
      // P = exp2(S * scale - m), bf16, into chunk `half` of the swizzled K-major tile (the previous P V has finished reading it)
      const float2 nm2 = make_float2(-m, -m);
      float2 sum = make_float2(0.f, 0.f);
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        const int hc = half * 2 + c;
        if (kDiag && hc > lg) {
#pragma unroll
          for (int u = 0; u < 4; ++u) *reinterpret_cast<uint4*>(chunk + swz(row, c * 4 + u)) = make_uint4(0u, 0u, 0u, 0u);
          continue;
        }
        uint32_t r[32];
        tmem_ld32(tS + c * 32, r);
        tmem_ld_wait();
        const bool cut = kDiag && hc == lg;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          uint32_t pk[4];
#pragma unroll
          for (int i = 0; i < 8; i += 2) {
            const float2 x = ffma2(u2f2(r[u * 8 + i], r[u * 8 + i + 1]), sc2, nm2);
            float2 e = make_float2(ex2(x.x), ex2(x.y));
            if (kDiag) { if (cut && u * 8 + i > lane) e.x = 0.f; if (cut && u * 8 + i + 1 > lane) e.y = 0.f; }
            sum = fadd2(sum, e);
            pk[i >> 1] = pack_bf16(e.x, e.y);
          }
          *reinterpret_cast<uint4*>(chunk + swz(row, c * 4 + u)) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
        }
      }
      l += sum.x + sum.y;
      tc_fence_before();
      fence_async_smem();                                    // generic-proxy stores -> visible to the tensor core's async proxy
      mbar_arrive(&bars->p_ready);
      ATTN_FSTAMP(g, 3);
      ++g;
    };
    for (int n = 0, w = attn_item(0, cta, G); w < nitems; ++n, w = attn_item(n, cta, G)) {
      const int qi = nq - 1 - w / p.BH, bh = w % p.BH, b = bh / p.H, h = bh % p.H;
#pragma unroll
      for (int i = 0; i < 16; ++i) o[i] = make_float2(0.f, 0.f);
      m = -INFINITY; l = 0.f;
      for (int j = 0; j < qi; ++j) tile(std::false_type{}, j, n > 0 && j == 0);
      tile(std::true_type{}, qi, n > 0 && qi == 0);
      // last block's P V
      mbar_wait(&bars->pv_full, (g - 1) & 1);
      tc_fence_after();
      {
        uint32_t r[32];
        tmem_ld32(tS0 + ((g - 1) & 1) * 128 + lane_addr + half * 32, r);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 16; ++i) o[i] = fadd2(o[i], u2f2(r[2 * i], r[2 * i + 1]));
      }
      tc_fence_before();
      if (half == 1) bars->xsum[row] = l;
      asm volatile("bar.sync 1, 256;" ::: "memory");
      if (half == 0) { l += bars->xsum[row]; bars->xsum[row] = l; }      // only this thread touches xsum[row] between the barriers
      asm volatile("bar.sync 1, 256;" ::: "memory");
      l = bars->xsum[row];
      const float inv = 1.f / l;
      // the output tile goes out through the copy engine (staged in the P tile, which the last P V has finished reading): a
      // thread's 64 bytes per row would otherwise be 32 separate transactions per store instruction
#pragma unroll
      for (int i = 0; i < 16; i += 4) {
        uint4 v = make_uint4(pack_bf16(o[i].x * inv, o[i].y * inv), pack_bf16(o[i + 1].x * inv, o[i + 1].y * inv), pack_bf16(o[i + 2].x * inv, o[i + 2].y * inv),
                             pack_bf16(o[i + 3].x * inv, o[i + 3].y * inv));
        *reinterpret_cast<uint4*>(p_smem + swz(row, half * 4 + (i >> 2))) = v;
      }
      fence_async_smem();
      asm volatile("bar.sync 1, 256;" ::: "memory");
      if (threadIdx.x == 0) { tma_store_2d(&map_out, p_smem, h * kD, b * p.T + qi * kBQ); bulk_commit(); }
      if (half == 0) p.lse[(size_t)bh * p.T + qi * kBQ + row] = m + log2f(l);
    }
    if (threadIdx.x == 0) bulk_wait_read0();
    ATTN_FSTAMP(30, 2);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 9) tmem_dealloc<256>(tmem);
  ATTN_FSTAMP(30, 3);
}

// =====================================================================================================
// Backward.  Work item = (128-row key/value block j, batch*head); query blocks i = j .. T/128 - 1 stream through it:
//   S  = Q_i K_j^T            dP = dO_i V_j^T                         (two 128x128x64 MMAs into TMEM)
//   P  = exp2(S * scale - lse_i)          dS = P o (dP - delta_i) * softmax_scale      (threads: row = query, written as bf16 into
//                                                                                         two swizzled [query][key] tiles)
//   dV += P^T dO_i            dK += dS^T Q_i            dQ_i = dS K_j  (three 128x64x128 MMAs; P / dS are read both MN-major - the
//                                                                        transposed products - and K-major from the same tiles)
// dV / dK accumulate in TMEM over the item's loop and are written once (bf16, straight into the dqkv matrix next to dq, i.e. in
// the layout the qkv projection's backward GEMMs read); dQ_i is added to an fp32 accumulator by the bulk-copy engine (several
// key blocks contribute to one query block) and converted by k_attn_dq_convert.
//
// PERSISTENT: one CTA per SM walks a list of items (heaviest first, the direction alternating from round to round so that
// every CTA gets a similar number of query blocks).  All three roles treat the query-block iterations of consecutive items
// as ONE stream (counter g): the producer fetches the next item's K / V (double-buffered) and Q / dO while the current item
// is still running, the MMA warp issues the next item's first S / dP ahead of the current item's last products, and the
// threads write the finished item's dK / dV while those run - so no per-item prologue or epilogue is exposed.
//   warp 8 TMA        warp 9 MMA issuer        warps 0-7: two threads per query row
// =====================================================================================================
struct __align__(16) AttnBwdBars {
  uint64_t kv_full[2], kv_empty[2], qdo_full[2], qdo_empty[2], sdp_full, sdp_taken, pds_ready, dq_full;
  uint32_t tmem_base;
};

struct AttnBwdParams {
  int T, H, C, BH;
  float scale_log2, scale;
  const float* lse;            // [B*H, T] log2 domain
  const float* delta;          // [B*H, T] rowsum(dO o O)
  float* dq_acc;               // fp32, zeroed by the caller: one 128 x 64 tile per (batch*head, query block), [16-byte unit][row]
  __nv_bfloat16* dqkv;         // [B*T, 3C]: dk / dv written here (columns C + h*64, 2C + h*64)
  long long* trace;            // debugging: clock64 stamps of CTA trace_x's first iterations (dcp_attn_set_trace), else nullptr
  int trace_x, trace_y;
};
#define ATTN_STAMP(g, k) do { if (p.trace != nullptr && (g) < 32 && blockIdx.x == p.trace_x && lane == 0 && (warp == 0 || warp == 9)) p.trace[(g) * 8 + (k)] = clock64(); } while (0)

// the r-th item of CTA c out of G: rounds alternate direction (c, 2G-1-c, 2G+c, ...); items are ordered heaviest first
__device__ __forceinline__ int bwd_item(int r, int c, int G) { return r * G + ((r & 1) ? G - 1 - c : c); }

__global__ void __launch_bounds__(kAttnThreads, 1)
k_attn_bwd(const __grid_constant__ CUtensorMap map_qkv, const __grid_constant__ CUtensorMap map_do, const __grid_constant__ CUtensorMap map_dqkv,
           const AttnBwdParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = align_smem_1024(smem_raw);
  uint8_t* kv_smem = smem;                                  // 2 buffers x (K 16 KB + V 16 KB)
  uint8_t* qdo_smem = smem + 4 * kTileBytes;                // 2 stages x (Q 16 KB + dO 16 KB)
  uint8_t* p_smem = qdo_smem + 4 * kTileBytes;              // 32 KB: P  [query][key], two 64-key chunks
  uint8_t* ds_smem = p_smem + 2 * kTileBytes;               // 32 KB: dS [query][key]
  uint8_t* dq_smem = ds_smem + 2 * kTileBytes;              // 32 KB: dQ_i staging for the bulk reduce-add (fp32, [16-byte unit][row])
  AttnBwdBars* bars = reinterpret_cast<AttnBwdBars*>(dq_smem + 2 * kTileBytes);
  const int warp = warp_idx_uniform(), lane = threadIdx.x & 31;
  const int nq = p.T / kBQ, nitems = nq * p.BH, G = gridDim.x, cta = blockIdx.x;

  griddep_launch();
  if (warp == 8 && lane == 0) { tma_prefetch_desc(&map_qkv); tma_prefetch_desc(&map_do); }
  if (warp == 9) {
    if (lane == 0) {
      for (int i = 0; i < 2; ++i) {
        mbar_init(&bars->kv_full[i], 1); mbar_init(&bars->kv_empty[i], 1);
        mbar_init(&bars->qdo_full[i], 1); mbar_init(&bars->qdo_empty[i], 1);
      }
      mbar_init(&bars->sdp_full, 1); mbar_init(&bars->sdp_taken, 256); mbar_init(&bars->pds_ready, 256); mbar_init(&bars->dq_full, 1);
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
    // ---- TMA producer: K_j / V_j of item n into buffer n & 1, then its (Q_i, dO_i) pairs through the two stages
    int g = 0;
    for (int n = 0, w = bwd_item(0, cta, G); w < nitems; ++n, w = bwd_item(n, cta, G)) {
      const int kj = w / p.BH, bh = w % p.BH, b = bh / p.H, h = bh % p.H, nit = nq - kj;
      mbar_wait(&bars->kv_empty[n & 1], ((n >> 1) & 1) ^ 1);
      if (lane == 0) {
        uint8_t* kb = kv_smem + (n & 1) * 2 * kTileBytes;
        mbar_expect_tx(&bars->kv_full[n & 1], 2 * kTileBytes);
        tma_load_2d(kb, &map_qkv, &bars->kv_full[n & 1], p.C + h * kD, b * p.T + kj * kBK);
        tma_load_2d(kb + kTileBytes, &map_qkv, &bars->kv_full[n & 1], 2 * p.C + h * kD, b * p.T + kj * kBK);
      }
      __syncwarp();
      for (int it = 0; it < nit; ++it, ++g) {
        const int st = g & 1;
        mbar_wait(&bars->qdo_empty[st], ((g >> 1) & 1) ^ 1);
        if (lane == 0) {
          const int r0 = b * p.T + (kj + it) * kBQ;
          mbar_expect_tx(&bars->qdo_full[st], 2 * kTileBytes);
          tma_load_2d(qdo_smem + st * 2 * kTileBytes, &map_qkv, &bars->qdo_full[st], h * kD, r0);
          tma_load_2d(qdo_smem + st * 2 * kTileBytes + kTileBytes, &map_do, &bars->qdo_full[st], h * kD, r0);
        }
        __syncwarp();
      }
    }
  } else if (warp == 9) {
    // ---- MMA issuer
    const uint32_t idesc_s = make_idesc(false, false, false, kBQ, kBK);      // S / dP: both operands K-major, 128 x 128
    const uint32_t idesc_t = make_idesc(false, true, true, kBK, kD);         // dV / dK: A = P^T / dS^T (MN-major), B = dO / Q (MN-major)
    const uint32_t idesc_q = make_idesc(false, false, true, kBQ, kD);        // dQ: A = dS (K-major), B = K (MN-major)
    const uint32_t pa = smem_u32(p_smem), dsa = smem_u32(ds_smem);
    // S / dP of stream iteration g, which belongs to the item whose K / V sit in buffer `kvb`
    auto issue_sdp = [&](int g, int kvb) {
      const int st = g & 1;
      mbar_wait(&bars->qdo_full[st], (g >> 1) & 1);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t qa = smem_u32(qdo_smem + st * 2 * kTileBytes), doa = qa + kTileBytes;
        const uint32_t ka = smem_u32(kv_smem + kvb * 2 * kTileBytes), va = ka + kTileBytes;
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
    int g = 0;
    if (bwd_item(0, cta, G) < nitems) {
      mbar_wait(&bars->kv_full[0], 0);
      issue_sdp(0, 0);
    }
    for (int n = 0, w = bwd_item(0, cta, G); w < nitems; ++n, w = bwd_item(n, cta, G)) {
      const int nit = nq - w / p.BH;
      const bool more = bwd_item(n + 1, cta, G) < nitems;
      const uint32_t ka = smem_u32(kv_smem + (n & 1) * 2 * kTileBytes);
      for (int it = 0; it < nit; ++it, ++g) {
        const int st = g & 1;
        // sdp_taken(g): S / dP of g are in the threads' registers: the next scores (of this item or of the next one) are
        // issued now, behind the previous iteration's products, and are ready when the threads have stored P / dS of g
        mbar_wait(&bars->sdp_taken, g & 1);
        tc_fence_after();
        ATTN_STAMP(g, 4);
        if (it + 1 < nit) issue_sdp(g + 1, n & 1);
        else if (more) { mbar_wait(&bars->kv_full[(n + 1) & 1], ((n + 1) >> 1) & 1); issue_sdp(g + 1, (n + 1) & 1); }
        ATTN_STAMP(g, 5);
        mbar_wait(&bars->pds_ready, g & 1);                // P / dS of g are in shared memory; dQ(g-1) and - at an item boundary -
        tc_fence_after();                                  // the finished item's dK / dV have been read
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
          for (int k = 0; k < kBK / 16; ++k)               // K index = key: dS chunk (k / 4), 32 B per step; K_j rows 16 k ..
            umma<false>(tDQ, make_smem_desc(dsa + (k >> 2) * kTileBytes + (k & 3) * 32, 0, 1024), make_smem_desc(ka + k * 2048, kTileBytes, 1024), idesc_q, k > 0);
          umma_commit(&bars->qdo_empty[st]);
          if (it == nit - 1) umma_commit(&bars->kv_empty[n & 1]);
          umma_commit(&bars->dq_full);
        }
        __syncwarp();
        ATTN_STAMP(g, 6);
      }
    }
  } else {
    // two threads per query row: `half` selects the 64 key columns of S / dP (= chunk `half` of the P / dS tiles), the 32 dQ
    // columns, and at the end of an item which of dK / dV this thread writes
    const int half = warp >> 2, lg = warp & 3;
    const int row = lg * 32 + lane;
    const uint32_t lane_addr = (uint32_t)(lg * 32) << 16;
    // dQ of stream iteration gq (query-block tile `qtile` of the accumulator) -> fp32 accumulator.  Several key blocks add into
    // one query block, so this is a reduction in global memory - done by the bulk-copy engine (one 32 KB cp.reduce.async.bulk
    // add per tile out of a staging buffer) rather than by per-lane red instructions, which cost the load/store pipe ~1.3
    // cycles per lane and instruction.  The accumulator tile of (head, query block) is contiguous, stored as [16-byte unit
    // 0..15][row 0..127]: conflict-free to stage from "thread = row", and k_attn_dq_convert undoes it.
    auto take_dq = [&](int gq, size_t qtile) {
      mbar_wait(&bars->dq_full, gq & 1);                    // all five products of iteration gq are done: dQ is complete and the
      tc_fence_after();                                     // P / dS tiles are no longer read
      ATTN_STAMP(gq, 7);
      uint32_t r[32];
      tmem_ld32(tDQ + lane_addr + half * 32, r);
      tmem_ld_wait();
      if (threadIdx.x == 0) bulk_wait_read0();              // the previous tile has left the staging buffer
      asm volatile("bar.sync 1, 256;" ::: "memory");
#pragma unroll
      for (int u = 0; u < 8; ++u)
        *reinterpret_cast<uint4*>(dq_smem + (((half * 8 + u) * kBQ + row) << 4)) = make_uint4(r[4 * u], r[4 * u + 1], r[4 * u + 2], r[4 * u + 3]);
      fence_async_smem();
      asm volatile("bar.sync 1, 256;" ::: "memory");
      if (threadIdx.x == 0) {
        bulk_reduce_add_f32(p.dq_acc + qtile * (kBQ * kD), dq_smem, kBQ * kD * 4);
        bulk_commit();
      }
    };
    // dK (half 0) / dV (half 1) of a finished item: thread = key row.  Staged (swizzled, 16 KB each) in the P tile - free once
    // the item's last products are done, and rewritten only after take_dq's wait on the bulk group below - and stored by the
    // copy engine: a thread's 128 bytes per row would be 32 separate transactions per store instruction.
    auto write_dkv = [&](int w, int gq) {
      const int kj = w / p.BH, bh = w % p.BH, b = bh / p.H, h = bh % p.H;
      mbar_wait(&bars->dq_full, gq & 1);
      tc_fence_after();
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        uint32_t r[32];
        tmem_ld32((half ? tDV : tDK) + lane_addr + c * 32, r);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; i += 8)
          *reinterpret_cast<uint4*>(p_smem + half * kTileBytes + swz(row, c * 4 + (i >> 3))) =
              make_uint4(pack_bf16(__uint_as_float(r[i]), __uint_as_float(r[i + 1])), pack_bf16(__uint_as_float(r[i + 2]), __uint_as_float(r[i + 3])),
                         pack_bf16(__uint_as_float(r[i + 4]), __uint_as_float(r[i + 5])), pack_bf16(__uint_as_float(r[i + 6]), __uint_as_float(r[i + 7])));
      }
      fence_async_smem();
      asm volatile("bar.sync 1, 256;" ::: "memory");
      if (threadIdx.x == 0) {
        tma_store_2d(&map_dqkv, p_smem, p.C + h * kD, b * p.T + kj * kBK);
        tma_store_2d(&map_dqkv, p_smem + kTileBytes, 2 * p.C + h * kD, b * p.T + kj * kBK);
        bulk_commit();
      }
    };
    uint8_t* pc = p_smem + half * kTileBytes;
    uint8_t* dc = ds_smem + half * kTileBytes;
    const float2 sc2 = make_float2(p.scale_log2, p.scale_log2), ss2 = make_float2(p.scale, p.scale);
    int g = 0, w_prev = -1;
    size_t qtile_prev = 0;
    // stream iteration g = query block `it` of item w; kDiag: the block on the diagonal (it == 0), the only one that needs the mask
    auto tile = [&](auto dtag, int w, int it) {
      constexpr bool kDiag = decltype(dtag)::value;
      const int kj = w / p.BH, bh = w % p.BH;
      const size_t grow = (size_t)bh * p.T + (kj + it) * kBQ + row;
      const float lse = p.lse[grow], delta = p.delta[grow];
      const float2 nl2 = make_float2(-lse, -lse), nd2 = make_float2(-delta * p.scale, -delta * p.scale);
      // P / dS of this iteration are computed into registers while the tensor core still runs the previous iteration's
      // dV / dK / dQ products (the MMA warp issues S / dP of this iteration ahead of them)
      uint32_t pk[32], dk[32];
      mbar_wait(&bars->sdp_full, g & 1);
      tc_fence_after();
      ATTN_STAMP(g, 0);
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        // diagonal block: chunk hc = 2 half + c is above the diagonal for hc > lg (P = dS = 0, nothing to read) and cut by it
        // for hc == lg (column e visible to lane >= e) - warp-uniform, as in the forward kernel
        const int hc = half * 2 + c;
        if (kDiag && hc > lg) {
#pragma unroll
          for (int e = 0; e < 16; ++e) { pk[c * 16 + e] = 0u; dk[c * 16 + e] = 0u; }
          continue;
        }
        const bool cut = kDiag && hc == lg;
        uint32_t rs[32], rp[32];
        tmem_ld32(tS + lane_addr + half * 64 + c * 32, rs);
        tmem_ld32(tDP + lane_addr + half * 64 + c * 32, rp);
        tmem_ld_wait();
#pragma unroll
        for (int e = 0; e < 32; e += 2) {
          const float2 x = ffma2(u2f2(rs[e], rs[e + 1]), sc2, nl2);
          float2 pr = make_float2(ex2(x.x), ex2(x.y));
          if (kDiag) { if (cut && e > lane) pr.x = 0.f; if (cut && e + 1 > lane) pr.y = 0.f; }
          const float2 d = fmul2(pr, ffma2(u2f2(rp[e], rp[e + 1]), ss2, nd2));     // P o (dP - delta) * scale
          pk[c * 16 + (e >> 1)] = pack_bf16(pr.x, pr.y);
          dk[c * 16 + (e >> 1)] = pack_bf16(d.x, d.y);
        }
      }
      tc_fence_before();
      mbar_arrive(&bars->sdp_taken);
      ATTN_STAMP(g, 1);
      if (kDiag && w_prev >= 0) write_dkv(w_prev, g - 1);   // item boundary: once the previous item's last products are done
      if (g > 0) take_dq(g - 1, qtile_prev);                // (waits until the copy engine has read the tiles staged just above)
      ATTN_STAMP(g, 2);
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        *reinterpret_cast<uint4*>(pc + swz(row, u)) = make_uint4(pk[u * 4], pk[u * 4 + 1], pk[u * 4 + 2], pk[u * 4 + 3]);
        *reinterpret_cast<uint4*>(dc + swz(row, u)) = make_uint4(dk[u * 4], dk[u * 4 + 1], dk[u * 4 + 2], dk[u * 4 + 3]);
      }
      tc_fence_before();
      fence_async_smem();
      mbar_arrive(&bars->pds_ready);                        // also: S / dP, the previous dQ and a finished item's dK / dV have been read
      ATTN_STAMP(g, 3);
      qtile_prev = (size_t)bh * nq + kj + it;
      ++g;
    };
    for (int n = 0, w = bwd_item(0, cta, G); w < nitems; ++n, w = bwd_item(n, cta, G)) {
      const int nit = nq - w / p.BH;
      tile(std::true_type{}, w, 0);
      w_prev = w;
      for (int it = 1; it < nit; ++it) tile(std::false_type{}, w, it);
    }
    if (g > 0) {
      write_dkv(w_prev, g - 1);
      take_dq(g - 1, qtile_prev);
      if (threadIdx.x == 0) bulk_wait0();                   // the last tile's reduction has been performed
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 9) tmem_dealloc<512>(tmem);
}

// delta[bh, t] = sum_d dO[t, h, d] * O[t, h, d]: eight lanes per (row, head), 16 bytes of O and dO each (consecutive lanes read
// consecutive memory), three shuffles.  The same threads zero the dQ accumulator (as many floats as O has elements).
__global__ void __launch_bounds__(256) k_attn_delta(const __nv_bfloat16* __restrict__ o, const __nv_bfloat16* __restrict__ dout, float* __restrict__ delta,
                                                    float* __restrict__ dq_acc, long nchunk, int T, int H) {
  griddep_launch();
  griddep_wait();
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;      // 16-byte chunk of the [B*T, H*64] matrices; nchunk is a multiple of 8
  float acc = 0.f;
  if (i < nchunk) {
    const uint4 a = reinterpret_cast<const uint4*>(o)[i], d = reinterpret_cast<const uint4*>(dout)[i];
    const __nv_bfloat162* a2 = reinterpret_cast<const __nv_bfloat162*>(&a);
    const __nv_bfloat162* d2 = reinterpret_cast<const __nv_bfloat162*>(&d);
#pragma unroll
    for (int k = 0; k < 4; ++k) acc += __low2float(a2[k]) * __low2float(d2[k]) + __high2float(a2[k]) * __high2float(d2[k]);
    reinterpret_cast<float4*>(dq_acc)[2 * i] = make_float4(0.f, 0.f, 0.f, 0.f);
    reinterpret_cast<float4*>(dq_acc)[2 * i + 1] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  acc += __shfl_xor_sync(0xffffffffu, acc, 1);
  acc += __shfl_xor_sync(0xffffffffu, acc, 2);
  acc += __shfl_xor_sync(0xffffffffu, acc, 4);
  if (i < nchunk && (threadIdx.x & 7) == 0) {
    const long rh = i >> 3, r = rh / H; const int h = (int)(rh % H);
    delta[((r / T) * H + h) * T + r % T] = acc;
  }
}

// dq accumulator tiles (fp32, one [16-byte unit][row] tile of 128 x 64 per (batch*head, query block)) -> bf16 into columns
// [0, C) of dqkv [B*T, 3C].  One CTA per tile; thread = (row, 32-column half).
__global__ void __launch_bounds__(256) k_attn_dq_convert(const float* __restrict__ acc, __nv_bfloat16* __restrict__ dqkv, int T, int H) {
  griddep_launch();
  griddep_wait();
  const int nq = T / kBQ, tile = blockIdx.x, bh = tile / nq, qi = tile % nq, b = bh / H, h = bh % H;
  const int row = threadIdx.x & (kBQ - 1), half = threadIdx.x >> 7;
  const uint4* src = reinterpret_cast<const uint4*>(acc + (size_t)tile * (kBQ * kD));
  __nv_bfloat16* dst = dqkv + ((size_t)b * T + qi * kBQ + row) * (3 * H * kD) + h * kD + half * 32;
  uint4 v[8];
#pragma unroll
  for (int u = 0; u < 8; ++u) v[u] = src[(half * 8 + u) * kBQ + row];
#pragma unroll
  for (int u = 0; u < 8; u += 2) {
    const float4 x = *reinterpret_cast<const float4*>(&v[u]), y = *reinterpret_cast<const float4*>(&v[u + 1]);
    *reinterpret_cast<uint4*>(dst + u * 4) = make_uint4(pack_bf16(x.x, x.y), pack_bf16(x.z, x.w), pack_bf16(y.x, y.y), pack_bf16(y.z, y.w));
  }
}

}  // namespace

extern "C" {

static long long *g_attn_trace = nullptr, *g_attn_ftrace = nullptr;
static int g_attn_trace_x = 0, g_attn_trace_y = 0;
// Debugging aid: device buffer (>= 256 long long) that one CTA fills with clock64 stamps, eight per key / query block (see
// ATTN_STAMP / ATTN_FSTAMP): CTA bx of the backward kernel, or - which = 1 - CTA (bx, 0) of the forward kernel.  nullptr = off.
void dcp_attn_set_trace(void* buf, int bx, int which) {
  (which ? g_attn_ftrace : g_attn_trace) = static_cast<long long*>(buf);
  g_attn_trace_x = bx; g_attn_trace_y = 0;
}

// Causal self-attention forward for head dimension 64.  qkv: bf16 [B*T, 3*H*64] (the fused projection: q | k | v, heads inside
// each third); out: bf16 [B*T, H*64]; lse: fp32 [B*H, T], log2-domain log-sum-exp of the scaled scores (consumed by
// dcp_attn_bwd).  T must be a multiple of 128.  softmax scale = 1/sqrt(64).
int dcp_attn_fwd(const void* qkv, void* out, float* lse, int B, int T, int H, cudaStream_t s) {
  if (T % kBQ != 0 || B < 1 || H < 1) { dcp_set_error("attention: T=%d must be a multiple of %d", T, kBQ); return -1; }
  const int C = H * kD;
  CUtensorMap map, map_out;
  if (encode_bf16_2d(&map, qkv, (long)B * T, 3L * C, 3L * C, kD, kBQ)) return -1;
  if (encode_bf16_2d(&map_out, out, (long)B * T, C, C, kD, kBQ)) return -1;
  AttnFwdParams p{};
  p.T = T; p.H = H; p.C = C; p.BH = B * H; p.scale_log2 = 0.125f * 1.4426950408889634f;
  p.out = static_cast<__nv_bfloat16*>(out); p.lse = lse; p.trace = g_attn_ftrace; p.trace_x = g_attn_trace_x;
  const size_t smem = (size_t)kTileBytes * (1 + kRing + 2) + sizeof(AttnBars) + 1024;
  // per device (a process may drive several GPUs): the shared-memory opt-in of the kernel and the SM count
  static int sms_of[64] = {0};
  int dev = 0;
  DCP_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) { dcp_set_error("attention: device index %d out of range", dev); return -1; }
  if (sms_of[dev] == 0) {
    DCP_CUDA_CHECK(cudaFuncSetAttribute(k_attn_fwd, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    DCP_CUDA_CHECK(cudaDeviceGetAttribute(&sms_of[dev], cudaDevAttrMultiProcessorCount, dev));
  }
  const int sms = sms_of[dev];
  const int items = B * H * (T / kBQ);
  static int per_sm = -1;           // DCP_ATTN_FWD_CTAS_PER_SM: 0 = one CTA per item (no persistence); default two persistent CTAs per SM
  if (per_sm < 0) { const char* e = getenv("DCP_ATTN_FWD_CTAS_PER_SM"); per_sm = e ? atoi(e) : 2; }
  const int grid = (per_sm <= 0 || items < per_sm * sms) ? items : per_sm * sms;
  DCP_CUDA_CHECK(dcp_launch_pdl(k_attn_fwd, dim3(grid), dim3(kAttnThreads), smem, s, map, map_out, p));
  return 0;
}


// Backward of dcp_attn_fwd.  dout / out: bf16 [B*T, H*64]; lse from the forward; dqkv: bf16 [B*T, 3*H*64] receives dq | dk | dv
// in the layout of qkv (what the projection's backward GEMMs consume - no split / merge copies).  Scratch from the caller:
// delta (B*H*T floats) and dq_acc (B*T*H*64 floats; zeroed here by the delta kernel).
int dcp_attn_bwd(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv, float* delta, float* dq_acc, int B, int T, int H,
                 cudaStream_t s) {
  if (T % kBQ != 0 || B < 1 || H < 1) { dcp_set_error("attention: T=%d must be a multiple of %d", T, kBQ); return -1; }
  const int C = H * kD;
  const long BT = (long)B * T;
  CUtensorMap mq, md, mg;
  if (encode_bf16_2d(&mq, qkv, BT, 3L * C, 3L * C, kD, kBQ)) return -1;
  if (encode_bf16_2d(&md, dout, BT, C, C, kD, kBQ)) return -1;
  if (encode_bf16_2d(&mg, dqkv, BT, 3L * C, 3L * C, kD, kBQ)) return -1;
  const long nchunk = BT * C / 8;
  DCP_CUDA_CHECK(dcp_launch_pdl(k_attn_delta, dim3((unsigned)((nchunk + 255) / 256)), dim3(256), 0, s, static_cast<const __nv_bfloat16*>(out),
                                static_cast<const __nv_bfloat16*>(dout), delta, dq_acc, nchunk, T, H));
  AttnBwdParams p{};
  p.T = T; p.H = H; p.C = C; p.BH = B * H; p.scale = 0.125f; p.scale_log2 = 0.125f * 1.4426950408889634f;
  p.lse = lse; p.delta = delta; p.dq_acc = dq_acc; p.dqkv = static_cast<__nv_bfloat16*>(dqkv); p.trace = g_attn_trace; p.trace_x = g_attn_trace_x; p.trace_y = g_attn_trace_y;
  const size_t smem = (size_t)kTileBytes * 14 + sizeof(AttnBwdBars) + 1024;
  // per device (a process may drive several GPUs): the shared-memory opt-in of the kernel and the SM count
  static int sms_of[64] = {0};
  int dev = 0;
  DCP_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) { dcp_set_error("attention: device index %d out of range", dev); return -1; }
  if (sms_of[dev] == 0) {
    DCP_CUDA_CHECK(cudaFuncSetAttribute(k_attn_bwd, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    DCP_CUDA_CHECK(cudaDeviceGetAttribute(&sms_of[dev], cudaDevAttrMultiProcessorCount, dev));
  }
  const int sms = sms_of[dev];
  const int items = B * H * (T / kBK);
  DCP_CUDA_CHECK(dcp_launch_pdl(k_attn_bwd, dim3(items < sms ? items : sms), dim3(kAttnThreads), smem, s, mq, md, mg, p));
  DCP_CUDA_CHECK(dcp_launch_pdl(k_attn_dq_convert, dim3((unsigned)(B * H * (T / kBQ))), dim3(256), 0, s, static_cast<const float*>(dq_acc), static_cast<__nv_bfloat16*>(dqkv), T, H));
  return 0;
}

}  // extern "C"
