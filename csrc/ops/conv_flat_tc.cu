// "Flat-shift" implicit-GEMM convolution on tcgen05 for valid (pad = 0, stride 1) convolutions whose
// activations are kept in the INPUT geometry: a tensor [N, H, W, C] is viewed as a matrix [N*H*W, C] of flat
// pixels, and the output of tap (r, s) for flat pixel p needs flat pixel p + r*W + s.  Positions outside the
// valid Ho x Wo window are stored as zeros, which makes forward, data-gradient and weight-gradient all plain
// shifted GEMMs over the same flat index (K3 of the reference model and its backward, ref:main.py:34,62).
//
// Why: the generic kernel (gemm_tc.cu) re-reads the activation window once per filter tap (9x the L2->SMEM
// traffic, which is what bounds it).  Here a persistent CTA keeps ALL filter taps resident in shared memory,
// loads each 128-pixel tile plus its halo ONCE with a single TMA box, and issues the 9 x (C/32) x 4 MMAs of the
// tile from shifted shared-memory descriptors (the matrix-descriptor base_offset field carries the swizzle
// phase of a window that does not start on an 8-row boundary).  Two TMEM accumulators let the epilogue of
// tile t overlap the MMAs of tile t+1.
//
// warp 0: TMA producer | warp 1: MMA issuer + TMEM owner | warps 2-5: epilogue.
#include <cudaTypedefs.h>

#include <cstdlib>

#include "ops/tc_common.cuh"

struct FlatConvParams {
  long P;                 // flat pixels (rows of in / out)
  int Cin, Cout;          // Cin in {32, 64}; Cout in {32, 64}
  int taps;               // <= 9
  int off[9];             // row shift of each tap (signed)
  int min_off, span;      // min(off), max(off) - min(off)
  int W, HW, Hv, Wv;      // row pitch, pixels per image, valid output extent (mode 0)
  int mode;               // 0: out = valid ? relu(acc + bias) : 0      1: out = mask > 0 ? acc : 0
                          // 2: relu(acc + bias) -> 2x2 max-pool -> channel dropout, written as pooled[N][Cout][Hv/2][Wv/2] + argmax bytes
  float* out; const float* bias; const float* mask;
  int tiles;
  int tiles_per_img, tile_rows;   // mode 2: a tile = 4 image rows (tile_rows = 4 W flat pixels), tiles_per_img = Hv / 4; else 0 / kBM
  uint8_t* arg;                   // mode 2: position (0..3) of the maximum inside each 2x2 window
  float drop_p; unsigned long long seed, stream; const long long* step; int training;   // mode 2: Philox channel dropout (ops_common.cuh)
  int use_base_offset;
  long long* dbg;         // optional: per-CTA clock64 trace (CTA 0 only); -DDCP_TRACE builds only
};

namespace {
using namespace tc;
constexpr int kIssuers = 3;                                  // MMA-issuing warps (1, 10, 11): the filter taps are split between them
constexpr int kBM = 128, kThreads = 32 * (9 + kIssuers), kTmem = 512;   // warps: 0 TMA, 1/10/11 MMA issuers, 2-9 epilogue (4 lane groups x 2 column chunks)

__device__ __forceinline__ uint64_t desc_k128(uint32_t addr, int use_base_offset) {
  // K-major, 128-byte swizzle, SBO = 1024 B; optional base_offset = row phase of the start address inside its 1024-B atom
  uint64_t d = make_smem_desc(addr, 0, 1024, 2);
  if (use_base_offset) d |= (uint64_t)((addr >> 7) & 7) << 49;
  return d;
}

struct __align__(16) Bars {
  uint64_t w_full, full[4], empty[4], acc_full[2], acc_empty[2];
  uint32_t tmem_base;
  uint32_t adelta[72], wdelta[72];     // per-MMA start-address increments (16-byte units) of the A / W descriptors
  alignas(16) float bias[64];
  alignas(16) float stage_out[8][32][36];          // per-epilogue-warp transpose buffer: coalesced 128-byte row stores
                                                   // (mode 2: one [Cout][kPoolPitch] tile of activations shared by the 8 epilogue warps)
  float keep[64];                                  // mode 2: dropout factor of each channel of the tile's image
};
constexpr int kPoolPitch = 105;                    // 4 x 26 pixels + 1: consecutive lanes (pixels) hit consecutive banks
static_assert(64 * kPoolPitch <= 8 * 32 * 36, "pool tile must fit in the transpose buffers");
__device__ __forceinline__ long flat_tile_start(const FlatConvParams& p, long tile) {
  return p.tiles_per_img ? (tile / p.tiles_per_img) * (long)p.HW + (tile % p.tiles_per_img) * (long)p.tile_rows : tile * (long)128;
}

template <int kMode>
__global__ void __launch_bounds__(kThreads, 1)
k_conv_flat(const __grid_constant__ CUtensorMap map_in, const __grid_constant__ CUtensorMap map_w, const FlatConvParams p, int stages, int halo_rows) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = align_smem_1024(smem_raw);
  const int cch = p.Cin / 32;
  const int w_bytes = p.taps * cch * p.Cout * 128;          // all taps resident
  const int chunk_bytes = halo_rows * 128;                   // one channel chunk of a haloed tile (multiple of 1024)
  const int stage_bytes = cch * chunk_bytes;
  uint8_t* w_smem = smem;
  uint8_t* a_smem = smem + ((w_bytes + 1023) & ~1023);
  Bars* bars = reinterpret_cast<Bars*>(a_smem + stages * stage_bytes);
  const int warp = warp_idx_uniform(), lane = threadIdx.x & 31;

  griddep_launch();                    // the next kernel may start its prologue (see common.cuh)
  if (warp == 0 && lane == 0) { tma_prefetch_desc(&map_in); tma_prefetch_desc(&map_w); }
  if (warp == 1) {
    if (lane == 0) {
      mbar_init(&bars->w_full, 1);
      for (int i = 0; i < 4; ++i) { mbar_init(&bars->full[i], 1); mbar_init(&bars->empty[i], kIssuers); }
      for (int i = 0; i < 2; ++i) { mbar_init(&bars->acc_full[i], kIssuers); mbar_init(&bars->acc_empty[i], 4 * (p.Cout / 32)); }
      mbar_fence_init();
    }
    __syncwarp();
    tmem_alloc<kTmem>(&bars->tmem_base);
  } else if (warp == 2) {
    // descriptor increments: MMA m = (tap, channel chunk, k-step)
    for (int m = lane; m < p.taps * cch * 4; m += 32) {
      const int j = m & 3, cc = (m >> 2) % cch, t = (m >> 2) / cch;
      bars->adelta[m] = (uint32_t)(((p.off[t] - p.min_off) * 128 + cc * chunk_bytes + j * 32) >> 4);
      bars->wdelta[m] = (uint32_t)(((t * cch + cc) * (p.Cout * 128) + j * 32) >> 4);
    }
  }
  griddep_wait();                      // everything above touched only shared memory / TMEM / kernel parameters
  if (warp == 2)
    for (int c = lane; c < p.Cout; c += 32) bars->bias[c] = p.bias ? p.bias[c] : 0.f;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *((volatile uint32_t*)&bars->tmem_base);
#ifdef DCP_TRACE      // clock64 trace of CTA 0 (build with DCP_NVCC_EXTRA=-DDCP_TRACE); compiled out of the production library
  long long* dbg = (p.dbg && blockIdx.x == 0) ? p.dbg : nullptr;
#else
  constexpr long long* dbg = nullptr;
#endif
  if (dbg && threadIdx.x == 0) dbg[0] = clock64();
  const int my_tiles = (p.tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;

  if (warp == 0) {
    // ---- producer: weights once, then one haloed tile per stage.  The whole warp walks the loop and stays converged
    //      (it reaches the closing __syncthreads as one); lane 0 issues.
    if (lane == 0) {
      mbar_expect_tx(&bars->w_full, w_bytes);
      for (int t = 0; t < p.taps; ++t)
        for (int cc = 0; cc < cch; ++cc)
          tma_load_2d(w_smem + (t * cch + cc) * (p.Cout * 128), &map_w, &bars->w_full, t * p.Cin + cc * 32, 0);
    }
    __syncwarp();
    for (int i = 0; i < my_tiles; ++i) {
      const int st = i % stages, ph = (i / stages) & 1;
      const long p0 = flat_tile_start(p, blockIdx.x + (long)i * gridDim.x);
      mbar_wait(&bars->empty[st], ph ^ 1);
      if (lane == 0) {
        mbar_expect_tx(&bars->full[st], stage_bytes);
        for (int cc = 0; cc < cch; ++cc)
          tma_load_2d(a_smem + st * stage_bytes + cc * chunk_bytes, &map_in, &bars->full[st], cc * 32, (int)(p0 + p.min_off));
      }
      __syncwarp();
    }
  } else if (warp == 1 || warp >= 10) {
    // ---- two MMA issuers (the single-thread issue rate, ~120 cycles per 128x64x8 instruction, is what bounds this
    //      kernel): warp 1 takes the first half of the filter taps, warp 10 the rest, each into its own TMEM
    //      accumulator; the epilogue adds the two.  Each warp runs the (uniform) loop, one elected lane issues.
    const int half = warp == 1 ? 0 : warp - 9;                        // issuer index 0 .. kIssuers-1
    const int per = (p.taps + kIssuers - 1) / kIssuers;
    const int t_begin = min(p.taps, half * per), t_end = min(p.taps, t_begin + per);
    const uint32_t idesc = make_idesc(true, false, false, kBM, p.Cout);
    const uint32_t w0 = smem_u32(w_smem);
    mbar_wait(&bars->w_full, 0);
    if (dbg && lane == 0 && half == 0) dbg[1] = clock64();
    for (int i = 0; i < my_tiles; ++i) {
      const int st = i % stages, ph = (i / stages) & 1;
      const int ab = i & 1, aph = (i >> 1) & 1;
      mbar_wait(&bars->acc_empty[ab], aph ^ 1);     // epilogue drained this accumulator
      if (dbg && i < 8 && lane == 0 && half == 0) dbg[8 + i * 8 + 0] = clock64();
      mbar_wait(&bars->full[st], ph);
      if (dbg && i < 8 && lane == 0 && half == 0) dbg[8 + i * 8 + 1] = clock64();
      tc_fence_after();
      const uint32_t a0 = smem_u32(a_smem + st * stage_bytes);
      const uint32_t tacc = tmem + (ab * kIssuers + half) * p.Cout;
      if (elect_one()) {
        // descriptor = constant high word | (address >> 4): everything below is computed from kernel parameters
        constexpr uint64_t kDescHi = ((uint64_t)(1024 >> 4) << 32) | ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
        uint32_t acc = 0;
        for (int t = t_begin; t < t_end; ++t) {
          const uint32_t arow = a0 + (uint32_t)(p.off[t] - p.min_off) * 128;
          for (int cc = 0; cc < cch; ++cc) {
            const uint32_t aa = arow + cc * chunk_bytes, wa = w0 + (t * cch + cc) * (p.Cout * 128);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              umma<true>(tacc, kDescHi | (uint64_t)(((aa + j * 32) >> 4) & 0x3FFF), kDescHi | (uint64_t)(((wa + j * 32) >> 4) & 0x3FFF), idesc, acc);
              acc = 1;
            }
          }
        }
        umma_commit(&bars->empty[st]);
        umma_commit(&bars->acc_full[ab]);
      }
      __syncwarp();
      if (dbg && i < 8 && lane == 0 && half == 0) dbg[8 + i * 8 + 2] = clock64();
    }
  } else if (warp >= 2 && warp < 10 && (warp - 2) / 4 < p.Cout / 32) {
    // ---- epilogue: warp = (lane group lg, 32-column chunk) ----
    const int lg = warp & 3, row = lg * 32 + lane, c0 = ((warp - 2) >> 2) * 32;
    float (*so)[36] = bars->stage_out[warp - 2];
    float4 bias4[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) bias4[j] = *reinterpret_cast<const float4*>(&bars->bias[c0 + 4 * j]);
    for (int i = 0; i < my_tiles; ++i) {
      const int ab = i & 1, aph = (i >> 1) & 1;
      const long tile0 = flat_tile_start(p, blockIdx.x + (long)i * gridDim.x);
      float vm = 1.f;
      if (kMode == 0) {
        const int q = (int)((tile0 + row) % p.HW);
        vm = ((q / p.W) < p.Hv && (q % p.W) < p.Wv) ? 1.f : 0.f;
      }
      mbar_wait(&bars->acc_full[ab], aph);
      if (dbg && i < 8 && warp == 2 && lane == 0) dbg[8 + i * 8 + 3] = clock64();
      tc_fence_after();
      uint32_t r[32];
      tmem_ld32(tmem + ((uint32_t)(lg * 32) << 16) + (ab * kIssuers) * p.Cout + c0, r);
#pragma unroll
      for (int h = 1; h < kIssuers; ++h) {      // add the partial accumulators of the other issuers
        uint32_t r2[32];
        tmem_ld32(tmem + ((uint32_t)(lg * 32) << 16) + (ab * kIssuers + h) * p.Cout + c0, r2);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) r[j] = __float_as_uint(__uint_as_float(r[j]) + __uint_as_float(r2[j]));
      }
      tmem_ld_wait();
      // the accumulator is in registers: hand the TMEM buffer back before the (long) store phase
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bars->acc_empty[ab]);
      if (dbg && i < 8 && warp == 2 && lane == 0) dbg[8 + i * 8 + 5] = clock64();
      if (kMode == 2) {
        // ---- fused 2x2 max-pool + channel dropout: the tile is 4 image rows (2 pooled rows) of one image.
        // phase 1: every epilogue warp drops relu(acc + bias) of its 32 pixels x 32 channels into the shared [channel][pixel] tile
        float* pt = &bars->stage_out[0][0][0];
        const int e = (warp - 2) * 32 + lane;                 // 0 .. 255 over the 8 epilogue warps
        const int img = (int)(tile0 / p.HW);
        if (row < p.tile_rows) {
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const float b = reinterpret_cast<const float*>(bias4)[j];
            pt[(c0 + j) * kPoolPitch + row] = fmaxf(__uint_as_float(r[j]) + b, 0.f);
          }
        }
        if (e < p.Cout) {
          float k = 1.f;
          if (p.training && p.drop_p > 0.f) {
            const unsigned long long st = p.stream + (p.step ? (unsigned long long)(*p.step) * 0x10000ull : 0ull);
            k = philox_uniform(p.seed, st, (unsigned long long)(img * p.Cout + e)) >= p.drop_p ? 1.f / (1.f - p.drop_p) : 0.f;
          }
          bars->keep[e] = k;
        }
        asm volatile("bar.sync 1, 256;" ::: "memory");
        // phase 2: per channel the two pooled rows of the tile are 2 * (Wv / 2) CONTIGUOUS outputs in the NCHW-flattened layout fc1 reads
        const int pwn = p.Wv >> 1, per_c = 2 * pwn, hp = p.Hv >> 1;
        const int rb = (int)((tile0 % p.HW) / p.tile_rows);   // row block of this tile inside its image
        for (int it = e; it < p.Cout * per_c; it += 256) {
          const int c = it / per_c, q = it % per_c, pr = q / pwn, pw = q % pwn;
          const float* src = pt + c * kPoolPitch + (2 * pr) * p.W + 2 * pw;
          const float v0 = src[0], v1 = src[1], v2 = src[p.W], v3 = src[p.W + 1];
          float m = v0; int a = 0;
          if (v1 > m) { m = v1; a = 1; }
          if (v2 > m) { m = v2; a = 2; }
          if (v3 > m) { m = v3; a = 3; }
          const long o = (((long)img * p.Cout + c) * hp + (rb * 2 + pr)) * pwn + pw;
          p.out[o] = m * bars->keep[c];
          p.arg[o] = (uint8_t)a;
        }
        asm volatile("bar.sync 1, 256;" ::: "memory");        // the tile is reused by the next iteration
        continue;
      }
      // lane = row: bias / ReLU / validity, then transpose through shared memory
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float4 v = make_float4(__uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1]), __uint_as_float(r[4 * j + 2]), __uint_as_float(r[4 * j + 3]));
        if (kMode == 0) {
          v.x = fmaxf(v.x + bias4[j].x, 0.f) * vm; v.y = fmaxf(v.y + bias4[j].y, 0.f) * vm;
          v.z = fmaxf(v.z + bias4[j].z, 0.f) * vm; v.w = fmaxf(v.w + bias4[j].w, 0.f) * vm;
        }
        *reinterpret_cast<float4*>(&so[lane][4 * j]) = v;
      }
      __syncwarp();
      if (dbg && i < 8 && warp == 2 && lane == 0) dbg[8 + i * 8 + 6] = clock64();
      // 8 lanes cover one 128-byte output row, 4 rows per instruction: coalesced stores (and mask loads)
      const long row0 = tile0 + lg * 32;
      const int c4 = (lane & 7) * 4, rsub = lane >> 3;
      float4 mk[8];
      if (kMode == 1) {
#pragma unroll
        for (int it = 0; it < 8; ++it) {
          const long prow = row0 + it * 4 + rsub;
          mk[it] = prow < p.P ? *reinterpret_cast<const float4*>(p.mask + prow * p.Cout + c0 + c4) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
      }
#pragma unroll
      for (int it = 0; it < 8; ++it) {
        const int rr = it * 4 + rsub;
        const long prow = row0 + rr;
        float4 v = *reinterpret_cast<const float4*>(&so[rr][c4]);
        if (kMode == 1) { v.x = mk[it].x > 0.f ? v.x : 0.f; v.y = mk[it].y > 0.f ? v.y : 0.f; v.z = mk[it].z > 0.f ? v.z : 0.f; v.w = mk[it].w > 0.f ? v.w : 0.f; }
        if (prow < p.P) *reinterpret_cast<float4*>(p.out + prow * p.Cout + c0 + c4) = v;
      }
      __syncwarp();
      if (dbg && i < 8 && warp == 2 && lane == 0) { dbg[8 + i * 8 + 7] = clock64(); dbg[8 + i * 8 + 4] = clock64(); }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (dbg && threadIdx.x == 0) dbg[2] = clock64();
  if (warp == 1) tmem_dealloc<kTmem>(tmem);
}

PFN_cuTensorMapEncodeTiled get_encode2() {
  static PFN_cuTensorMapEncodeTiled fn = nullptr;
  if (!fn) {
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void**)&fn, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) fn = nullptr;
  }
  return fn;
}
int map2d(CUtensorMap* m, const void* base, long rows, long cols, int box_cols, int box_rows) {
  auto enc = get_encode2();
  if (!enc) { dcp_set_error("cuTensorMapEncodeTiled unavailable"); return -1; }
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)cols * 4};
  cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { dcp_set_error("flat conv: cuTensorMapEncodeTiled(rows=%ld cols=%ld box=%dx%d) failed: %d", rows, cols, box_cols, box_rows, (int)r); return -1; }
  return 0;
}
}  // namespace

extern "C" {

// in: [P, Cin] fp32 flat pixels (pitch W, HW pixels per image); wmat: [Cout][taps*Cin] fp32 (tap-major columns);
// off[t]: flat row shift of tap t.  mode 0: out = (h < Hv && w < Wv) ? relu(acc + bias) : 0;  mode 1: out = mask > 0 ? acc : 0.
static int conv_flat_launch(const float* in, const float* wmat, float* out, const float* bias, const float* mask, long P, int Cin, int Cout, int taps,
                            const int* off, int W, int HW, int Hv, int Wv, int mode, int num_ctas, const FlatConvParams* extra, cudaStream_t s) {
  if ((Cin != 32 && Cin != 64) || (Cout != 32 && Cout != 64) || taps < 1 || taps > 9) { dcp_set_error("flat conv: unsupported Cin/Cout/taps %d/%d/%d", Cin, Cout, taps); return -1; }
  if (taps < kIssuers) { dcp_set_error("flat conv: needs at least %d filter taps (one per issuing warp)", kIssuers); return -1; }
  FlatConvParams p{};
  p.P = P; p.Cin = Cin; p.Cout = Cout; p.taps = taps; p.W = W; p.HW = HW; p.Hv = Hv; p.Wv = Wv; p.mode = mode;
  p.out = out; p.bias = bias; p.mask = mask;
  int mn = off[0], mx = off[0];
  for (int t = 0; t < taps; ++t) { p.off[t] = off[t]; mn = off[t] < mn ? off[t] : mn; mx = off[t] > mx ? off[t] : mx; }
  p.min_off = mn; p.span = mx - mn;
  p.tiles = (int)((P + kBM - 1) / kBM);
  p.tile_rows = kBM;
  if (mode == 2) {
    // a tile = 4 image rows: the 2x2 windows of two pooled rows live inside one accumulator tile
    if ((Hv & 3) || (Wv & 1) || 4 * W > kBM || Cout != 64 || P % HW) { dcp_set_error("flat conv + pool: needs Hv %% 4 == 0, even Wv, 4 * W <= %d, Cout == 64", kBM); return -1; }
    p.tiles_per_img = Hv / 4; p.tile_rows = 4 * W; p.tiles = (int)(P / HW) * p.tiles_per_img;
    p.arg = extra->arg; p.drop_p = extra->drop_p; p.seed = extra->seed; p.stream = extra->stream; p.step = extra->step; p.training = extra->training;
  }
  { static const int base_off = [] { const char* e = getenv("DCP_FLAT_BASEOFF"); return e ? atoi(e) : 0; }(); p.use_base_offset = base_off; }
#ifdef DCP_TRACE
  { const char* e = getenv("DCP_FLAT_DBG_PTR"); p.dbg = e ? (long long*)strtoull(e, nullptr, 0) : nullptr; }
#else
  p.dbg = nullptr;
#endif
  const int halo_rows = (kBM + p.span + 7) / 8 * 8;
  if (halo_rows > 256) { dcp_set_error("flat conv: halo of %d rows exceeds one TMA box", halo_rows); return -1; }
  const int cch = Cin / 32;
  const int w_bytes = ((taps * cch * Cout * 128) + 1023) & ~1023;
  const int stage_bytes = cch * halo_rows * 128;
  int stages = 4;
  const int budget = 222 * 1024;
  while (stages > 1 && w_bytes + stages * stage_bytes + (int)sizeof(Bars) + 1024 > budget) --stages;
  const size_t smem = (size_t)w_bytes + (size_t)stages * stage_bytes + sizeof(Bars) + 1024;
  if (smem > (size_t)budget) { dcp_set_error("flat conv: needs %zu bytes of shared memory", smem); return -1; }
  static bool attr = false;
  if (!attr) {
    DCP_CUDA_CHECK(cudaFuncSetAttribute(k_conv_flat<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, budget));
    DCP_CUDA_CHECK(cudaFuncSetAttribute(k_conv_flat<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, budget));
    DCP_CUDA_CHECK(cudaFuncSetAttribute(k_conv_flat<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, budget));
    attr = true;
  }
  CUtensorMap mi, mw;
  if (map2d(&mi, in, P, Cin, 32, halo_rows)) return -1;
  if (map2d(&mw, wmat, Cout, (long)taps * Cin, 32, Cout)) return -1;
  if (num_ctas <= 0) num_ctas = 148;
  if (num_ctas > p.tiles) num_ctas = p.tiles;
  if (mode == 0) DCP_CUDA_CHECK(dcp_launch_pdl(k_conv_flat<0>, dim3(num_ctas), dim3(kThreads), smem, s, mi, mw, p, stages, halo_rows));
  else if (mode == 1) DCP_CUDA_CHECK(dcp_launch_pdl(k_conv_flat<1>, dim3(num_ctas), dim3(kThreads), smem, s, mi, mw, p, stages, halo_rows));
  else DCP_CUDA_CHECK(dcp_launch_pdl(k_conv_flat<2>, dim3(num_ctas), dim3(kThreads), smem, s, mi, mw, p, stages, halo_rows));
  DCP_CUDA_CHECK(cudaGetLastError());
  return 0;
}

int dcp_conv_flat_tc(const float* in, const float* wmat, float* out, const float* bias, const float* mask, long P, int Cin, int Cout, int taps,
                     const int* off, int W, int HW, int Hv, int Wv, int mode, int num_ctas, cudaStream_t s) {
  if (mode != 0 && mode != 1) { dcp_set_error("flat conv: mode must be 0 or 1 (the pooling epilogue is dcp_conv_flat_pool_tc)"); return -1; }
  return conv_flat_launch(in, wmat, out, bias, mask, P, Cin, Cout, taps, off, W, HW, Hv, Wv, mode, num_ctas, nullptr, s);
}

// Convolution + bias + ReLU + 2x2 max-pool + channel dropout in ONE kernel (K3-K6 of the reference model, ref:main.py:34-37): the
// conv output never reaches memory.  pooled: [N][Cout][Hv/2][Wv/2] fp32 (the NCHW-flattened order fc1 consumes, ref:main.py:38-39),
// arg: one byte per pooled element = position of the maximum in its window (for the backward pass).  Dropout drops whole channels
// of an image with probability drop_p (scaled by 1/(1-p)), Philox stream `stream` + step counter, regenerated - not stored - in
// backward.  Needs Hv % 4 == 0, Wv even, 4 * W <= 128 flat pixels, Cout == 64.
int dcp_conv_flat_pool_tc(const float* in, const float* wmat, float* pooled, void* arg, const float* bias, long P, int Cin, int Cout, int taps,
                          const int* off, int W, int HW, int Hv, int Wv, float drop_p, unsigned long long seed, unsigned long long stream,
                          const void* step, int training, int num_ctas, cudaStream_t s) {
  FlatConvParams e{};
  e.arg = static_cast<uint8_t*>(arg); e.drop_p = drop_p; e.seed = seed; e.stream = stream; e.step = static_cast<const long long*>(step); e.training = training;
  return conv_flat_launch(in, wmat, pooled, bias, nullptr, P, Cin, Cout, taps, off, W, HW, Hv, Wv, 2, num_ctas, &e, s);
}

}  // extern "C"
