// Persistent bf16 GEMM for sm_100a - the Linear layers of the transformer / ResNet configurations (forward, data-gradient and
// weight-gradient through the operand-major flags, K8 / K12 of the reference model generalised; ref:main.py:39,43).
//
//   C[M,N] = op(A) . op(B)^T   fp32 accumulation in TMEM, output bf16 or fp32, optional bias + ReLU
//
// Why a second GEMM kernel next to gemm_tc.cu: that one launches one CTA per 128 x 128 tile and drains the accumulator
// after the main loop - prologue, epilogue and main loop are serial, and with N = 128 every tcgen05.mma is worth ~64 tensor
// cycles while one thread can issue one only every ~128 cycles.  Here
//   * the tile is 128 x 256 (one MMA = 128 x 256 x 16 = 128 tensor cycles: the issue rate and the tensor pipe balance),
//   * ONE CTA per SM walks the tile list (persistent): barrier init, TMEM allocation and descriptor prefetch happen once,
//   * the accumulator is double buffered in TMEM (2 x 256 columns): eight epilogue warps drain tile i (tcgen05.ld ->
//     bias / ReLU / cast -> shared-memory transpose -> coalesced stores) while the MMA warp already runs tile i + 1,
//   * a 4-stage TMA pipeline of 48 KB stages (A 128 x 64, B 256 x 64 bf16, 128-byte swizzle) feeds it,
//   * split-K work units (tile, k-slice) write fp32 partial tiles for dcp_splitk_reduce when a problem has too few tiles.
// Warp roles: 0 = TMA producer, 1 = MMA issuer (+ TMEM allocator), 2..9 = epilogue (lane group = warp % 4, column half = (warp-2)/4).
#include <cstdlib>

#include "ops/tc_common.cuh"

struct BfGemmParams {
  int M, N;
  int k_stages;             // K / 64
  int BN;                   // 128 or 256
  int a_mn, b_mn;           // operand majors (0: K-major [rows][K]; 1: MN-major [K][rows])
  int tiles_m, tiles_n, ksplit;
  int stages;               // pipeline depth
  void* out; long ldc; int out_bf16;
  const void* bias; int bias_bf16; int relu;
  long partial_stride;      // ksplit > 1: slice z writes fp32 to out + z * partial_stride
  // conv = 1: implicit-GEMM convolution (stride 1 or 2).  An M tile is a box of Nb x Hb x Wb = 128 output pixels, K stage ks is
  // (filter tap, 64-channel chunk): A = 4-D TMA box of the channels-last activation shifted by the tap (OOB zero fill = padding,
  // elementStrides = convolution stride), B = the matching 64 columns of the [Kout][R*S*C] filter matrix.
  int conv, S, cchunks, Cin, pad, cstride, Wb, Hb, Nb, tiles_w, tiles_h, Ho, Wo, Nimg;
  // Tail split (ksplit == 1 only): tiles [0, tail_t0) are whole tiles; each tile of the last, partial wave [tail_t0, tiles) is cut
  // into tail_ks K slices that other (otherwise idle) CTAs take, writing tile-local fp32 partials [128][BN] to tail_ws; the
  // fix-up kernel adds them up.  148 + 44 tiles then cost 1 + 1/3 waves instead of 2.
  int tail_t0, tail_ks;
  float* tail_ws;
};

namespace {
using namespace tc;

constexpr int kBM = 128, kBK = 64, kMaxStages = 6;
constexpr int kThreads = 320, kEpiWarps = 8;
constexpr int kABytes = kBM * 128;                       // 16 KB
constexpr int kStgPitch = 20;                            // floats per staged row (16 columns + pad: conflict-free 16-byte accesses)
constexpr int kStgBytes = kEpiWarps * 32 * kStgPitch * 4;

struct __align__(8) PPipe {
  uint64_t full[kMaxStages], empty[kMaxStages], acc_full[2], acc_empty[2];
  uint32_t tmem_base;
};

struct Unit { int m_tile, n_tile, ks0, ks1, z; float* part; };
template <bool kTail> __device__ __forceinline__ Unit decode_unit(const BfGemmParams& p, int u, int tiles, int kper) {
  Unit w;
  if (!kTail || p.ksplit > 1 || u < p.tail_t0) {
    w.z = u / tiles; const int t = u % tiles;
    w.m_tile = t % p.tiles_m; w.n_tile = t / p.tiles_m;
    w.ks0 = w.z * kper; w.ks1 = min(p.k_stages, w.ks0 + kper); w.part = nullptr;
  } else {
    const int r = u - p.tail_t0, t = p.tail_t0 + r / p.tail_ks, sl = r % p.tail_ks;
    const int kp = (p.k_stages + p.tail_ks - 1) / p.tail_ks;
    w.z = 0; w.m_tile = t % p.tiles_m; w.n_tile = t / p.tiles_m;
    w.ks0 = sl * kp; w.ks1 = min(p.k_stages, w.ks0 + kp);
    w.part = p.tail_ws + (size_t)r * kBM * p.BN;
  }
  return w;
}

// kTail: the launch uses the tail split (a separate instantiation keeps the common path free of its bookkeeping)
template <bool kTail>
__global__ void __launch_bounds__(kThreads, 1)
k_gemm_bf16_persistent(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b, const BfGemmParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = align_smem_1024(smem_raw);
  const int b_bytes = p.BN * 128, stage_bytes = kABytes + b_bytes, nstages = p.stages;
  float* stg_all = reinterpret_cast<float*>(smem + nstages * stage_bytes);
  PPipe* pipe = reinterpret_cast<PPipe*>(smem + nstages * stage_bytes + kStgBytes);
  const int warp = warp_idx_uniform(), lane = threadIdx.x & 31;

  griddep_launch();
  if (warp == 0 && lane == 0) { tma_prefetch_desc(&map_a); tma_prefetch_desc(&map_b); }
  if (warp == 1) {
    if (lane == 0) {
      for (int i = 0; i < kMaxStages; ++i) { mbar_init(&pipe->full[i], 1); mbar_init(&pipe->empty[i], 1); }
      for (int i = 0; i < 2; ++i) { mbar_init(&pipe->acc_full[i], 1); mbar_init(&pipe->acc_empty[i], kEpiWarps); }
      mbar_fence_init();
    }
    __syncwarp();
    tmem_alloc<512>(&pipe->tmem_base);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *((volatile uint32_t*)&pipe->tmem_base);
  griddep_wait();

  // work units: u -> (tile, k-slice); tiles run M-fastest so that one wave of CTAs shares a few B tiles
  const int tiles = p.tiles_m * p.tiles_n;
  const int units = (!kTail || p.ksplit > 1) ? tiles * p.ksplit : p.tail_t0 + (tiles - p.tail_t0) * p.tail_ks;
  const int kper = (p.k_stages + p.ksplit - 1) / p.ksplit;

  if (warp == 0) {
    // ================= TMA producer (the whole warp walks the loop, one lane issues: the warp stays converged and
    // reaches the closing __syncthreads as one) =================
    int it = 0;
    for (int u = blockIdx.x; u < units; u += gridDim.x) {
      const Unit w = decode_unit<kTail>(p, u, tiles, kper);
      const int m_tile = w.m_tile, n_tile = w.n_tile, ks0 = w.ks0, ks1 = w.ks1;
      for (int ks = ks0; ks < ks1; ++ks, ++it) {
        const int st = it % nstages, ph = (it / nstages) & 1;
        mbar_wait(&pipe->empty[st], ph ^ 1);
        if (lane == 0) {
          uint8_t* sa = smem + st * stage_bytes;
          uint8_t* sb = sa + kABytes;
          mbar_expect_tx(&pipe->full[st], stage_bytes);
          int kb = ks * kBK;
          if (p.conv) {
            int tt = m_tile;
            const int tw = tt % p.tiles_w; tt /= p.tiles_w;
            const int th = tt % p.tiles_h; tt /= p.tiles_h;
            const int tap = ks / p.cchunks, cc = ks % p.cchunks;
            tma_load_4d(sa, &map_a, &pipe->full[st], cc * kBK, tw * p.Wb * p.cstride - p.pad + tap % p.S, th * p.Hb * p.cstride - p.pad + tap / p.S, tt * p.Nb);
            kb = tap * p.Cin + cc * kBK;
          } else if (!p.a_mn) tma_load_2d(sa, &map_a, &pipe->full[st], kb, m_tile * kBM);
          else
            for (int j = 0; j < kBM / 64; ++j) tma_load_2d(sa + j * (kBK * 128), &map_a, &pipe->full[st], m_tile * kBM + j * 64, kb);
          if (!p.b_mn) tma_load_2d(sb, &map_b, &pipe->full[st], kb, n_tile * p.BN);
          else
            for (int j = 0; j < p.BN / 64; ++j) tma_load_2d(sb + j * (kBK * 128), &map_b, &pipe->full[st], n_tile * p.BN + j * 64, kb);
        }
        __syncwarp();
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    const uint32_t idesc = make_idesc(false, p.a_mn != 0, p.b_mn != 0, kBM, p.BN);
    const uint64_t step_a = p.a_mn ? (uint64_t)((16 * 128) >> 4) : 2u, step_b = p.b_mn ? (uint64_t)((16 * 128) >> 4) : 2u;
    int it = 0, tcount = 0;
    for (int u = blockIdx.x; u < units; u += gridDim.x, ++tcount) {
      const Unit w = decode_unit<kTail>(p, u, tiles, kper);
      const int ks0 = w.ks0, ks1 = w.ks1;
      const int ab = tcount & 1, aph = (tcount >> 1) & 1;
      mbar_wait(&pipe->acc_empty[ab], aph ^ 1);            // epilogue has drained this accumulator buffer
      tc_fence_after();
      const uint32_t tacc = tmem + ab * 256;
      for (int ks = ks0; ks < ks1; ++ks, ++it) {
        const int st = it % nstages, ph = (it / nstages) & 1;
        mbar_wait(&pipe->full[st], ph);
        tc_fence_after();
        const uint32_t sa = smem_u32(smem + st * stage_bytes), sb = sa + kABytes;
        if (elect_one()) {
          const uint64_t da0 = p.a_mn ? make_smem_desc(sa, kBK * 128, 1024) : make_smem_desc(sa, 0, 1024);
          const uint64_t db0 = p.b_mn ? make_smem_desc(sb, kBK * 128, 1024) : make_smem_desc(sb, 0, 1024);
#pragma unroll
          for (int j = 0; j < kBK / 16; ++j) umma<false>(tacc, da0 + j * step_a, db0 + j * step_b, idesc, (ks > ks0 || j > 0) ? 1u : 0u);
          umma_commit(&pipe->empty[st]);
          if (ks == ks1 - 1) umma_commit(&pipe->acc_full[ab]);
        }
        __syncwarp();
      }
      if (ks1 <= ks0 && elect_one()) mbar_arrive(&pipe->acc_full[ab]);   // empty k-slice: nothing to wait for
      __syncwarp();
    }
  } else {
    // ================= epilogue: TMEM -> registers -> smem transpose -> coalesced global stores =================
    const int ew = warp - 2, lg = warp & 3, chalf = ew >> 2;
    float* stg = stg_all + ew * (32 * kStgPitch);
    const int cpw = p.BN / 2, cbeg = chalf * cpw;
    int tcount = 0;
    for (int u = blockIdx.x; u < units; u += gridDim.x, ++tcount) {
      const Unit w = decode_unit<kTail>(p, u, tiles, kper);
      const int z = w.z, m_tile = w.m_tile, n_tile = w.n_tile, ks0 = w.ks0, ks1 = w.ks1;
      const bool have = ks1 > ks0;
      const int ab = tcount & 1, aph = (tcount >> 1) & 1;
      mbar_wait(&pipe->acc_full[ab], aph);
      tc_fence_after();
      const bool add_bias = p.bias != nullptr && p.partial_stride == 0 && w.part == nullptr;
      const long zoff = p.partial_stride ? (long)z * p.partial_stride : 0;
      const long row0 = (long)m_tile * kBM + lg * 32;
      long orow[4];                                            // output row of the 4 tile rows this lane stores (-1: outside)
#pragma unroll
      for (int itr = 0; itr < 4; ++itr) {
        const int rt = lg * 32 + itr * 8 + (lane >> 2);         // row inside the 128-row tile
        if (p.conv) {
          int tt = m_tile;
          const int tw = tt % p.tiles_w; tt /= p.tiles_w;
          const int th = tt % p.tiles_h; tt /= p.tiles_h;
          const int n = tt * p.Nb + rt / (p.Hb * p.Wb), h = th * p.Hb + (rt / p.Wb) % p.Hb, w = tw * p.Wb + rt % p.Wb;
          orow[itr] = (n < p.Nimg && h < p.Ho && w < p.Wo) ? ((long)n * p.Ho + h) * p.Wo + w : -1;
        } else {
          orow[itr] = (long)m_tile * kBM + rt < p.M ? (long)m_tile * kBM + rt : -1;
        }
      }
      (void)row0;
      if (p.out_bf16 && have && w.part == nullptr) {
        // bf16 output (the common case, and the one whose epilogue must keep up with a 12-stage main loop): 32 columns per
        // TMEM load, converted to bf16 BEFORE the shared-memory transpose (half the staging traffic), 64 contiguous bytes
        // per row and store; the TMEM load of the next chunk is in flight while this one is stored.
        uint32_t* stw = reinterpret_cast<uint32_t*>(stg);                 // [32 rows][20 words]: 16 words of data + pad
        uint32_t r[32];
        tmem_ld32(tmem + ((uint32_t)(lg * 32) << 16) + ab * 256 + cbeg, r);
        for (int c0 = cbeg; c0 < cbeg + cpw; c0 += 32) {
          tmem_ld_wait();
          const int gn0 = n_tile * p.BN + c0;
          const int ncol = min(32, p.N - gn0);
          uint32_t w[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            float v0 = __uint_as_float(r[2 * j]), v1 = __uint_as_float(r[2 * j + 1]);
            if (add_bias) {
              const int g0 = min(gn0 + 2 * j, p.N - 1), g1 = min(gn0 + 2 * j + 1, p.N - 1);
              v0 += p.bias_bf16 ? __bfloat162float(((const __nv_bfloat16*)p.bias)[g0]) : ((const float*)p.bias)[g0];
              v1 += p.bias_bf16 ? __bfloat162float(((const __nv_bfloat16*)p.bias)[g1]) : ((const float*)p.bias)[g1];
            }
            if (p.relu) { v0 = fmaxf(v0, 0.f); v1 = fmaxf(v1, 0.f); }
            const __nv_bfloat162 t = __floats2bfloat162_rn(v0, v1);
            w[j] = *reinterpret_cast<const uint32_t*>(&t);
          }
          if (c0 + 32 < cbeg + cpw) tmem_ld32(tmem + ((uint32_t)(lg * 32) << 16) + ab * 256 + c0 + 32, r);   // next chunk, asynchronous
          else {                                               // that was the last read of this buffer: hand it back to the MMA warp
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&pipe->acc_empty[ab]);
          }
          if (ncol > 0) {
#pragma unroll
            for (int j = 0; j < 4; ++j) *reinterpret_cast<uint4*>(stw + lane * kStgPitch + j * 4) = make_uint4(w[4 * j], w[4 * j + 1], w[4 * j + 2], w[4 * j + 3]);
          }
          __syncwarp();
          const int c8 = (lane & 3) * 8, nv = min(8, ncol - c8);      // 8 bf16 = 16 bytes per lane, 4 lanes per row
          if (ncol > 0) {
#pragma unroll
            for (int itr = 0; itr < 4; ++itr) {
              const int rl = itr * 8 + (lane >> 2);
              const long grow = orow[itr];
              if (grow < 0 || nv <= 0) continue;
              const uint4 v = *reinterpret_cast<const uint4*>(stw + rl * kStgPitch + (lane & 3) * 4);
              __nv_bfloat16* o = (__nv_bfloat16*)p.out + grow * p.ldc + gn0 + c8;
              if (nv == 8 && (reinterpret_cast<uintptr_t>(o) & 15) == 0) *reinterpret_cast<uint4*>(o) = v;
              else {
                const __nv_bfloat16* ve = reinterpret_cast<const __nv_bfloat16*>(&v);
                for (int e = 0; e < nv; ++e) o[e] = ve[e];
              }
            }
          }
          __syncwarp();
        }
        continue;
      }
      for (int c0 = cbeg; c0 < cbeg + cpw; c0 += 16) {
        uint32_t r[16];
        if (have) { tmem_ld16(tmem + ((uint32_t)(lg * 32) << 16) + ab * 256 + c0, r); tmem_ld_wait(); }
        else {
#pragma unroll
          for (int j = 0; j < 16; ++j) r[j] = 0;
        }
        if (c0 + 16 >= cbeg + cpw) {                         // last read of this buffer: hand it back to the MMA warp
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&pipe->acc_empty[ab]);
        }
        const int gn0 = n_tile * p.BN + c0;
        const bool relu = p.relu && w.part == nullptr;
        const int ncol = w.part ? 16 : min(16, p.N - gn0);
        if (ncol <= 0) continue;
        if (add_bias || relu) {
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const int gn = gn0 + j;
            float v = __uint_as_float(r[j]);
            if (add_bias && gn < p.N) v += p.bias_bf16 ? __bfloat162float(((const __nv_bfloat16*)p.bias)[gn]) : ((const float*)p.bias)[gn];
            if (relu) v = fmaxf(v, 0.f);
            r[j] = __float_as_uint(v);
          }
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) *reinterpret_cast<uint4*>(stg + lane * kStgPitch + j * 4) = make_uint4(r[4 * j], r[4 * j + 1], r[4 * j + 2], r[4 * j + 3]);
        __syncwarp();
        const int c4 = (lane & 3) * 4, nv = min(4, ncol - c4);
#pragma unroll
        for (int itr = 0; itr < 4; ++itr) {
          const int rl = itr * 8 + (lane >> 2);
          if (w.part) {                                      // tail slice: tile-local fp32 partial, every row and column of the tile
            *reinterpret_cast<float4*>(w.part + (size_t)(lg * 32 + rl) * p.BN + c0 + c4) = *reinterpret_cast<const float4*>(stg + rl * kStgPitch + c4);
            continue;
          }
          const long grow = orow[itr];
          if (grow < 0 || nv <= 0) continue;
          const float4 v = *reinterpret_cast<const float4*>(stg + rl * kStgPitch + c4);
          if (p.out_bf16) {
            __nv_bfloat16* o = (__nv_bfloat16*)p.out + grow * p.ldc + gn0 + c4;
            if (nv == 4 && (reinterpret_cast<uintptr_t>(o) & 7) == 0) {
              const __nv_bfloat162 t0 = __floats2bfloat162_rn(v.x, v.y), t1 = __floats2bfloat162_rn(v.z, v.w);
              *reinterpret_cast<uint2*>(o) = make_uint2(*reinterpret_cast<const uint32_t*>(&t0), *reinterpret_cast<const uint32_t*>(&t1));
            } else {
              const float ve[4] = {v.x, v.y, v.z, v.w};
              for (int e = 0; e < nv; ++e) o[e] = __float2bfloat16_rn(ve[e]);
            }
          } else {
            float* o = (float*)p.out + zoff + grow * p.ldc + gn0 + c4;
            if (nv == 4 && (reinterpret_cast<uintptr_t>(o) & 15) == 0) *reinterpret_cast<float4*>(o) = v;
            else {
              const float ve[4] = {v.x, v.y, v.z, v.w};
              for (int e = 0; e < nv; ++e) o[e] = ve[e];
            }
          }
        }
        __syncwarp();
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<512>(tmem);
}

// Fix-up of the tail split: out tile = sum of its tail_ks tile-local partials (+ bias, ReLU, cast).  One thread = 8 columns of a row.
__global__ void __launch_bounds__(256) k_gemm_tail_fixup(const BfGemmParams p) {
  griddep_launch();
  griddep_wait();
  const int tiles = p.tiles_m * p.tiles_n, ntail = tiles - p.tail_t0;
  const int vpr = p.BN / 8;                                    // 8-column vectors per tile row
  const long total = (long)ntail * kBM * vpr;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int v = (int)(i % vpr); long r_ = i / vpr; const int row = (int)(r_ % kBM); const int tt = (int)(r_ / kBM);
    const int t = p.tail_t0 + tt, m_tile = t % p.tiles_m, n_tile = t / p.tiles_m;
    const long grow = (long)m_tile * kBM + row;
    const int gn0 = n_tile * p.BN + v * 8;
    if (grow >= p.M || gn0 >= p.N) continue;
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
    for (int sl = 0; sl < p.tail_ks; ++sl) {
      const float* src = p.tail_ws + ((size_t)(tt * p.tail_ks + sl) * kBM + row) * p.BN + v * 8;
      const float4 a = *reinterpret_cast<const float4*>(src), b = *reinterpret_cast<const float4*>(src + 4);
      acc[0] += a.x; acc[1] += a.y; acc[2] += a.z; acc[3] += a.w; acc[4] += b.x; acc[5] += b.y; acc[6] += b.z; acc[7] += b.w;
    }
    const int nv = min(8, p.N - gn0);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if (p.bias && j < nv) acc[j] += p.bias_bf16 ? __bfloat162float(((const __nv_bfloat16*)p.bias)[gn0 + j]) : ((const float*)p.bias)[gn0 + j];
      if (p.relu) acc[j] = fmaxf(acc[j], 0.f);
    }
    if (p.out_bf16) {
      __nv_bfloat16* o = (__nv_bfloat16*)p.out + grow * p.ldc + gn0;
      if (nv == 8 && (reinterpret_cast<uintptr_t>(o) & 15) == 0) {
        uint4 pk; __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
        for (int j = 0; j < 4; ++j) h[j] = __floats2bfloat162_rn(acc[2 * j], acc[2 * j + 1]);
        *reinterpret_cast<uint4*>(o) = pk;
      } else for (int j = 0; j < nv; ++j) o[j] = __float2bfloat16_rn(acc[j]);
    } else {
      float* o = (float*)p.out + grow * p.ldc + gn0;
      for (int j = 0; j < nv; ++j) o[j] = acc[j];
    }
  }
}

// workspace for the tail split, registered by the host binding (the library itself does not allocate): one per process / device
float* g_tail_ws = nullptr;
size_t g_tail_ws_bytes = 0;

}  // namespace

extern "C" {

// Hand the GEMM a scratch buffer (device memory, 16-byte aligned) for the tail split of two-wave problems; nullptr / 0 disables it.
// GEMMs that use it must not overlap on different streams.
int dcp_gemm_set_workspace(void* ws, size_t bytes) { g_tail_ws = static_cast<float*>(ws); g_tail_ws_bytes = bytes; return 0; }

// bf16 operands; see gemm_tc.cu::dcp_tc_gemm_ex for the meaning of the major flags.  ksplit > 1: `out` is a
// [ksplit][M * ldc] fp32 scratch that receives partial tiles (finish with dcp_splitk_reduce, which also adds the bias).
// K must be a multiple of 64; MN-major operands need their row count (M resp. N) to be a multiple of 8 (TMA stride rule).
int dcp_gemm_bf16_persistent(const void* A, const void* B, void* out, const void* bias, int M, int N, int K, long lda, long ldb, long ldc,
                             int a_mn, int b_mn, int out_bf16, int bias_bf16, int relu, int ksplit, int num_ctas, cudaStream_t s) {
  if (K % kBK) { dcp_set_error("bf16 gemm: K=%d must be a multiple of %d", K, kBK); return -1; }
  if (ksplit < 1) ksplit = 1;
  if (ksplit > 1 && (out_bf16 || relu)) { dcp_set_error("bf16 gemm: split-K writes fp32 partials (no bf16 output / ReLU)"); return -1; }
  BfGemmParams p{};
  p.M = M; p.N = N; p.k_stages = K / kBK; p.a_mn = a_mn; p.b_mn = b_mn; p.ksplit = ksplit;
  p.BN = N > 128 ? 256 : 128;
  p.tiles_m = (M + kBM - 1) / kBM; p.tiles_n = (N + p.BN - 1) / p.BN;
  // few work units even after the caller's K split: halve the tile so that more SMs get work
  if (p.BN == 256 && (long)p.tiles_m * p.tiles_n * ksplit < 64) { p.BN = 128; p.tiles_n = (N + 127) / 128; }
  p.out = out; p.ldc = ldc; p.out_bf16 = out_bf16; p.bias = bias; p.bias_bf16 = bias_bf16; p.relu = relu;
  p.partial_stride = ksplit > 1 ? (long)M * ldc : 0;
  const int stage_bytes = kABytes + p.BN * 128;
  const int budget = 227 * 1024 - kStgBytes - (int)sizeof(PPipe) - 1024;
  p.stages = budget / stage_bytes;
  if (p.stages > kMaxStages) p.stages = kMaxStages;
  if (p.stages < 2) { dcp_set_error("bf16 gemm: stage does not fit shared memory"); return -1; }
  const size_t smem = (size_t)p.stages * stage_bytes + kStgBytes + sizeof(PPipe) + 1024;
  static bool attr = false;
  if (!attr) {
    DCP_CUDA_CHECK(cudaFuncSetAttribute(k_gemm_bf16_persistent<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    DCP_CUDA_CHECK(cudaFuncSetAttribute(k_gemm_bf16_persistent<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    attr = true;
  }
  CUtensorMap ma, mb;
  if (!a_mn) { if (encode_bf16_2d(&ma, A, M, K, lda, kBK, kBM)) return -1; }
  else { if (encode_bf16_2d(&ma, A, K, M, lda, 64, kBK)) return -1; }
  if (!b_mn) { if (encode_bf16_2d(&mb, B, N, K, ldb, kBK, p.BN)) return -1; }
  else { if (encode_bf16_2d(&mb, B, K, N, ldb, 64, kBK)) return -1; }
  if (num_ctas <= 0) num_ctas = 148;
  const int tiles = p.tiles_m * p.tiles_n;
  p.tail_t0 = tiles; p.tail_ks = 1; p.tail_ws = nullptr;
  // tail split: the last wave holds few tiles and K is long enough to cut -> spread those tiles over the idle CTAs
  const int tail = tiles % num_ctas;
  if (ksplit == 1 && tiles > num_ctas && tail > 0 && tail <= num_ctas / 2 && p.k_stages >= 16 && g_tail_ws != nullptr) {
    int ks = num_ctas / tail;
    if (ks > p.k_stages / 8) ks = p.k_stages / 8;
    if (ks > 8) ks = 8;
    const size_t need = (size_t)tail * ks * kBM * p.BN * sizeof(float);
    if (ks >= 2 && need <= g_tail_ws_bytes) { p.tail_t0 = tiles - tail; p.tail_ks = ks; p.tail_ws = g_tail_ws; }
  }
  const int units = ksplit > 1 ? tiles * ksplit : p.tail_t0 + (tiles - p.tail_t0) * p.tail_ks;
  if (num_ctas > units) num_ctas = units;
  if (p.tail_ws) DCP_CUDA_CHECK(dcp_launch_pdl(k_gemm_bf16_persistent<true>, dim3(num_ctas), dim3(kThreads), smem, s, ma, mb, p));
  else DCP_CUDA_CHECK(dcp_launch_pdl(k_gemm_bf16_persistent<false>, dim3(num_ctas), dim3(kThreads), smem, s, ma, mb, p));
  if (p.tail_ws) {
    const long total = (long)(tiles - p.tail_t0) * kBM * (p.BN / 8);
    long g = (total + 255) / 256; if (g > 148 * 8) g = 148 * 8;
    DCP_CUDA_CHECK(dcp_launch_pdl(k_gemm_tail_fixup, dim3((unsigned)g), dim3(256), 0, s, p));
  }
  return 0;
}


// Implicit-GEMM convolution on the persistent kernel (bf16, stride 1 or 2): y[n,ho,wo,k] = sum x[n, ho*stride-pad+r, wo*stride-pad+s, c] * w[k][(r*S+s)*C + c]
//   x: channels-last [N,H,W,C] (C a multiple of 64); wmat: [Kout][R*S*C] row-major; y: channels-last [N,Ho,Wo,Kout] bf16 or fp32.
//   (Wb, Hb, Nb): box of output pixels per M tile, product 128.  The data-gradient is the same call on dy with flipped filters.
int dcp_conv_bf16_persistent(const void* x, const void* wmat, void* y, const void* bias, int N, int H, int W, int C, int Kout, int R, int S, int pad,
                             int stride, int Ho, int Wo, int Wb, int Hb, int Nb, int out_bf16, int bias_bf16, int relu, cudaStream_t s) {
  if (C % kBK) { dcp_set_error("bf16 conv: C=%d must be a multiple of %d", C, kBK); return -1; }
  if (Wb * Hb * Nb != kBM) { dcp_set_error("bf16 conv: box %dx%dx%d must cover %d pixels", Wb, Hb, Nb, kBM); return -1; }
  if (stride < 1 || stride > 2 || Wb * stride > 256 || Hb * stride > 256) { dcp_set_error("bf16 conv: stride %d / box %dx%d not supported", stride, Wb, Hb); return -1; }
  BfGemmParams p{};
  p.conv = 1; p.S = S; p.cchunks = C / kBK; p.Cin = C; p.pad = pad; p.cstride = stride; p.Wb = Wb; p.Hb = Hb; p.Nb = Nb;
  p.tiles_w = (Wo + Wb - 1) / Wb; p.tiles_h = (Ho + Hb - 1) / Hb; p.Ho = Ho; p.Wo = Wo; p.Nimg = N;
  p.tiles_m = p.tiles_w * p.tiles_h * ((N + Nb - 1) / Nb);
  p.M = p.tiles_m * kBM; p.N = Kout; p.k_stages = R * S * p.cchunks; p.ksplit = 1;
  p.BN = Kout > 128 ? 256 : 128;
  p.tiles_n = (Kout + p.BN - 1) / p.BN;
  if (p.BN == 256 && (long)p.tiles_m * p.tiles_n < 64) { p.BN = 128; p.tiles_n = (Kout + 127) / 128; }
  p.out = y; p.ldc = Kout; p.out_bf16 = out_bf16; p.bias = bias; p.bias_bf16 = bias_bf16; p.relu = relu;
  p.tail_t0 = p.tiles_m * p.tiles_n; p.tail_ks = 1; p.tail_ws = nullptr;      // no tail split in convolution mode
  const int stage_bytes = kABytes + p.BN * 128;
  const int budget = 227 * 1024 - kStgBytes - (int)sizeof(PPipe) - 1024;
  p.stages = budget / stage_bytes;
  if (p.stages > kMaxStages) p.stages = kMaxStages;
  const size_t smem = (size_t)p.stages * stage_bytes + kStgBytes + sizeof(PPipe) + 1024;
  static bool attr = false;
  if (!attr) {
    DCP_CUDA_CHECK(cudaFuncSetAttribute(k_gemm_bf16_persistent<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    DCP_CUDA_CHECK(cudaFuncSetAttribute(k_gemm_bf16_persistent<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    attr = true;
  }
  CUtensorMap ma, mb;
  if (encode_bf16_4d(&ma, x, N, H, W, C, kBK, Wb, Hb, Nb, stride)) return -1;
  if (encode_bf16_2d(&mb, wmat, Kout, (long)R * S * C, (long)R * S * C, kBK, p.BN)) return -1;
  int num_ctas = p.tiles_m * p.tiles_n;
  if (num_ctas > 148) num_ctas = 148;
  DCP_CUDA_CHECK(dcp_launch_pdl(k_gemm_bf16_persistent<false>, dim3(num_ctas), dim3(kThreads), smem, s, ma, mb, p));
  return 0;
}

}  // extern "C"
