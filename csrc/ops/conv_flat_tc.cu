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
  float* out; const float* bias; const float* mask;
  int tiles;
  int use_base_offset;
};

namespace {
using namespace tc;
constexpr int kBM = 128, kThreads = 192, kTmem = 128;

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
  alignas(16) float stage_out[4][32][36];          // per-epilogue-warp transpose buffer: coalesced 128-byte row stores
};

__global__ void __launch_bounds__(kThreads, 1)
k_conv_flat(const __grid_constant__ CUtensorMap map_in, const __grid_constant__ CUtensorMap map_w, const FlatConvParams p, int stages, int halo_rows) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int cch = p.Cin / 32;
  const int w_bytes = p.taps * cch * p.Cout * 128;          // all taps resident
  const int chunk_bytes = halo_rows * 128;                   // one channel chunk of a haloed tile (multiple of 1024)
  const int stage_bytes = cch * chunk_bytes;
  uint8_t* w_smem = smem;
  uint8_t* a_smem = smem + ((w_bytes + 1023) & ~1023);
  Bars* bars = reinterpret_cast<Bars*>(a_smem + stages * stage_bytes);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) { tma_prefetch_desc(&map_in); tma_prefetch_desc(&map_w); }
  if (warp == 1) {
    if (lane == 0) {
      mbar_init(&bars->w_full, 1);
      for (int i = 0; i < 4; ++i) { mbar_init(&bars->full[i], 1); mbar_init(&bars->empty[i], 1); }
      for (int i = 0; i < 2; ++i) { mbar_init(&bars->acc_full[i], 1); mbar_init(&bars->acc_empty[i], 4); }
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
    for (int c = lane; c < p.Cout; c += 32) bars->bias[c] = p.bias ? p.bias[c] : 0.f;
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *((volatile uint32_t*)&bars->tmem_base);
  const int my_tiles = (p.tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;

  if (warp == 0 && lane == 0) {
    // ---- producer: weights once, then one haloed tile per stage ----
    mbar_expect_tx(&bars->w_full, w_bytes);
    for (int t = 0; t < p.taps; ++t)
      for (int cc = 0; cc < cch; ++cc)
        tma_load_2d(w_smem + (t * cch + cc) * (p.Cout * 128), &map_w, &bars->w_full, t * p.Cin + cc * 32, 0);
    for (int i = 0; i < my_tiles; ++i) {
      const int st = i % stages, ph = (i / stages) & 1;
      const long p0 = (long)(blockIdx.x + (long)i * gridDim.x) * kBM;
      mbar_wait(&bars->empty[st], ph ^ 1);
      mbar_expect_tx(&bars->full[st], stage_bytes);
      for (int cc = 0; cc < cch; ++cc)
        tma_load_2d(a_smem + st * stage_bytes + cc * chunk_bytes, &map_in, &bars->full[st], cc * 32, (int)(p0 + p.min_off));
    }
  } else if (warp == 1 && lane == 0) {
    // ---- MMA issuer ----
    const uint32_t idesc = make_idesc(true, false, false, kBM, p.Cout);
    mbar_wait(&bars->w_full, 0);
    for (int i = 0; i < my_tiles; ++i) {
      const int st = i % stages, ph = (i / stages) & 1;
      const int ab = i & 1, aph = (i >> 1) & 1;
      mbar_wait(&bars->acc_empty[ab], aph ^ 1);     // epilogue drained this accumulator
      mbar_wait(&bars->full[st], ph);
      tc_fence_after();
      const uint64_t a_base = desc_k128(smem_u32(a_smem + st * stage_bytes), 0), w_base = desc_k128(smem_u32(w_smem), 0);
      const uint32_t tacc = tmem + ab * p.Cout;
      const int nmma = p.taps * cch * 4;
      // the start-address field (16-byte units, < 2^14) is the only part of the descriptor that changes
      umma<true>(tacc, a_base + bars->adelta[0], w_base + bars->wdelta[0], idesc, 0);
#pragma unroll 4
      for (int m = 1; m < nmma; ++m) umma<true>(tacc, a_base + bars->adelta[m], w_base + bars->wdelta[m], idesc, 1);
      umma_commit(&bars->empty[st]);
      umma_commit(&bars->acc_full[ab]);
    }
  } else if (warp >= 2) {
    // ---- epilogue ----
    const int lg = warp & 3, row = lg * 32 + lane;
    for (int i = 0; i < my_tiles; ++i) {
      const int ab = i & 1, aph = (i >> 1) & 1;
      const long pp = (long)(blockIdx.x + (long)i * gridDim.x) * kBM + row;
      mbar_wait(&bars->acc_full[ab], aph);
      tc_fence_after();
      bool valid = true;
      if (p.mode == 0) {
        const int q = (int)(pp % p.HW);
        valid = (q / p.W) < p.Hv && (q % p.W) < p.Wv;
      }
      float (*so)[36] = bars->stage_out[warp - 2];
      const long tile_row0 = (long)(blockIdx.x + (long)i * gridDim.x) * kBM + lg * 32;   // first row this warp owns
      for (int c0 = 0; c0 < p.Cout; c0 += 32) {
        uint32_t r[32];
        tmem_ld32(tmem + ((uint32_t)(lg * 32) << 16) + ab * p.Cout + c0, r);
        tmem_ld_wait();
        // lane = row: bias / ReLU / validity here, then transpose through shared memory
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
          float4 v = make_float4(__uint_as_float(r[j]), __uint_as_float(r[j + 1]), __uint_as_float(r[j + 2]), __uint_as_float(r[j + 3]));
          if (p.mode == 0) {
            const float4 b = *reinterpret_cast<const float4*>(&bars->bias[c0 + j]);
            v.x = valid ? fmaxf(v.x + b.x, 0.f) : 0.f; v.y = valid ? fmaxf(v.y + b.y, 0.f) : 0.f;
            v.z = valid ? fmaxf(v.z + b.z, 0.f) : 0.f; v.w = valid ? fmaxf(v.w + b.w, 0.f) : 0.f;
          }
          *reinterpret_cast<float4*>(&so[lane][j]) = v;
        }
        __syncwarp();
        // 8 lanes cover one 128-byte output row, 4 rows per instruction: fully coalesced stores (and mask loads)
#pragma unroll
        for (int it = 0; it < 8; ++it) {
          const int rr = it * 4 + (lane >> 3), c4 = (lane & 7) * 4;
          const long prow = tile_row0 + rr;
          if (prow < p.P) {
            float4 v = *reinterpret_cast<const float4*>(&so[rr][c4]);
            if (p.mode == 1) {
              const float4 mk = *reinterpret_cast<const float4*>(p.mask + prow * p.Cout + c0 + c4);
              v.x = mk.x > 0.f ? v.x : 0.f; v.y = mk.y > 0.f ? v.y : 0.f; v.z = mk.z > 0.f ? v.z : 0.f; v.w = mk.w > 0.f ? v.w : 0.f;
            }
            *reinterpret_cast<float4*>(p.out + prow * p.Cout + c0 + c4) = v;
          }
        }
        __syncwarp();
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bars->acc_empty[ab]);   // 4 epilogue warps -> count 4
    }
  }
  tc_fence_before();
  __syncthreads();
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
int dcp_conv_flat_tc(const float* in, const float* wmat, float* out, const float* bias, const float* mask, long P, int Cin, int Cout, int taps,
                     const int* off, int W, int HW, int Hv, int Wv, int mode, int num_ctas, cudaStream_t s) {
  if ((Cin != 32 && Cin != 64) || (Cout != 32 && Cout != 64) || taps < 1 || taps > 9) { dcp_set_error("flat conv: unsupported Cin/Cout/taps %d/%d/%d", Cin, Cout, taps); return -1; }
  FlatConvParams p{};
  p.P = P; p.Cin = Cin; p.Cout = Cout; p.taps = taps; p.W = W; p.HW = HW; p.Hv = Hv; p.Wv = Wv; p.mode = mode;
  p.out = out; p.bias = bias; p.mask = mask;
  int mn = off[0], mx = off[0];
  for (int t = 0; t < taps; ++t) { p.off[t] = off[t]; mn = off[t] < mn ? off[t] : mn; mx = off[t] > mx ? off[t] : mx; }
  p.min_off = mn; p.span = mx - mn;
  p.tiles = (int)((P + kBM - 1) / kBM);
  { const char* e = getenv("DCP_FLAT_BASEOFF"); p.use_base_offset = e ? atoi(e) : 0; }
  const int halo_rows = (kBM + p.span + 7) / 8 * 8;
  if (halo_rows > 256) { dcp_set_error("flat conv: halo of %d rows exceeds one TMA box", halo_rows); return -1; }
  const int cch = Cin / 32;
  const int w_bytes = ((taps * cch * Cout * 128) + 1023) & ~1023;
  const int stage_bytes = cch * halo_rows * 128;
  int stages = 4;
  const int budget = 200 * 1024;
  while (stages > 1 && w_bytes + stages * stage_bytes + (int)sizeof(Bars) + 1024 > budget) --stages;
  const size_t smem = (size_t)w_bytes + (size_t)stages * stage_bytes + sizeof(Bars) + 1024;
  if (smem > (size_t)budget) { dcp_set_error("flat conv: needs %zu bytes of shared memory", smem); return -1; }
  static bool attr = false;
  if (!attr) { DCP_CUDA_CHECK(cudaFuncSetAttribute(k_conv_flat, cudaFuncAttributeMaxDynamicSharedMemorySize, budget)); attr = true; }
  CUtensorMap mi, mw;
  if (map2d(&mi, in, P, Cin, 32, halo_rows)) return -1;
  if (map2d(&mw, wmat, Cout, (long)taps * Cin, 32, Cout)) return -1;
  if (num_ctas <= 0) num_ctas = 148;
  if (num_ctas > p.tiles) num_ctas = p.tiles;
  k_conv_flat<<<num_ctas, kThreads, smem, s>>>(mi, mw, p, stages, halo_rows);
  DCP_CUDA_CHECK(cudaGetLastError());
  return 0;
}

}  // extern "C"
