// Flat-shift convolution WEIGHT gradient on tcgen05 (companion of conv_flat_tc.cu; backward of K3, ref:main.py:34,62):
//   dw[k][r][s][c] = sum over flat pixels p of  dy[p][k] * x[p + r*W + s][c]
// with x = [P, 32] flat pixels (pitch W) and dy = [P, Cout] in the same frame (zeros outside the valid window, so
// out-of-window positions contribute nothing).  Pixels run along the MMA K axis; both operands are read MN-major
// straight from the activation tensors (tf32 => SWIZZLE_128B_BASE32B layout, TMA 128B_ATOM_32B swizzle).
//
// One persistent CTA owns a contiguous range of 128-pixel stages.  Per stage ONE TMA box brings the x tile plus halo
// (128 + 2W + 3 rows) and two boxes bring the dy tile.  Filter row r has its own issuing warp and TMEM accumulator
// (128 lanes = 4 chunks of 32 channels, chunk s = tap (r, s); chunk 3 is padding): its A operand is the x tile
// shifted by r*W rows whose 128-byte M-chunks are ONE row apart (LBO = 128 B) - the three horizontal taps are
// overlapping windows of the same shared-memory tile, so x is loaded once instead of 9 times.
// While the tensor core works, the four epilogue warps add up the columns of the dy tile in shared memory: that is the
// BIAS gradient db[k] = sum_p dy[p][k], which so costs no extra pass over dy.
// Partial filters and bias gradients of the CTAs go to a scratch buffer [CTA][Cout*288 + Cout] and are summed by
// dcp_splitk_reduce2.
#include <cudaTypedefs.h>

#include "ops/tc_common.cuh"

struct WgradFlatParams {
  long P;
  int Cout, W;              // Cout in {32, 64}; row pitch of the frame
  int stages_total;         // ceil(P / 128)
  float* scratch;           // [gridDim.x][Cout*288 + Cout]
};

namespace {
using namespace tc;
constexpr int kBK = 128;                 // pixels per stage
constexpr int kThreads = 32 * 8;         // warp 0 TMA, warps 1-3 MMA issuers (filter rows), warps 4-7 epilogue
constexpr int kTmem = 256;               // 3 accumulators x 64 columns

struct __align__(16) WBars {
  uint64_t full[3], empty[3], acc_full;
  uint32_t tmem_base;
};

// MN-major tf32 descriptor: BASE32B layout (type 1), SBO = 4 k-rows = 512 B, LBO = distance between 128-byte M/N chunks
__device__ __forceinline__ uint64_t desc_mn32(uint32_t addr, uint32_t lbo_bytes) { return make_smem_desc(addr, lbo_bytes, 512, 1); }

__global__ void __launch_bounds__(kThreads, 1)
k_wgrad_flat(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_dy, const WgradFlatParams p, int nstages, int halo_rows) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = align_smem_1024(smem_raw);
  const int x_bytes = halo_rows * 128;                     // multiple of 1024
  const int dy_chunks = p.Cout / 32, dy_bytes = dy_chunks * kBK * 128;
  const int stage_bytes = x_bytes + dy_bytes;
  WBars* bars = reinterpret_cast<WBars*>(smem + nstages * stage_bytes);
  const int warp = warp_idx_uniform(), lane = threadIdx.x & 31;

  griddep_launch();
  if (warp == 0 && lane == 0) { tma_prefetch_desc(&map_x); tma_prefetch_desc(&map_dy); }
  if (warp == 1) {
    if (lane == 0) {
      for (int i = 0; i < 3; ++i) { mbar_init(&bars->full[i], 1); mbar_init(&bars->empty[i], 3 + 4); }
      mbar_init(&bars->acc_full, 3);
      mbar_fence_init();
    }
    __syncwarp();
    tmem_alloc<kTmem>(&bars->tmem_base);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  griddep_wait();                      // prologue done; from here on global memory of earlier kernels is read
  const uint32_t tmem = *((volatile uint32_t*)&bars->tmem_base);
  // contiguous range of stages for this CTA
  const int per = (p.stages_total + gridDim.x - 1) / gridDim.x;
  const int s0 = blockIdx.x * per, s1 = min(p.stages_total, s0 + per);
  const int n = max(0, s1 - s0);

  if (warp == 0) {
    // producer: the whole warp walks the loop (stays converged up to the closing __syncthreads), lane 0 issues
    for (int i = 0; i < n; ++i) {
      const int st = i % nstages, ph = (i / nstages) & 1;
      const int row0 = (s0 + i) * kBK;
      uint8_t* sx = smem + st * stage_bytes;
      mbar_wait(&bars->empty[st], ph ^ 1);
      if (lane == 0) {
        mbar_expect_tx(&bars->full[st], stage_bytes);
        tma_load_2d(sx, &map_x, &bars->full[st], 0, row0);
        for (int j = 0; j < dy_chunks; ++j) tma_load_2d(sx + x_bytes + j * (kBK * 128), &map_dy, &bars->full[st], j * 32, row0);
      }
      __syncwarp();
    }
  } else if (warp >= 1 && warp <= 3) {
    // issuer for filter row r: A = x tile shifted by r*W rows, M-chunks one row apart; B = dy tile
    const int r = warp - 1;
    const uint32_t idesc = make_idesc(true, true, true, 128, p.Cout);
    const uint32_t tacc = tmem + r * 64;
    for (int i = 0; i < n; ++i) {
      const int st = i % nstages, ph = (i / nstages) & 1;
      mbar_wait(&bars->full[st], ph);
      tc_fence_after();
      const uint32_t sx = smem_u32(smem + st * stage_bytes), sdy = sx + x_bytes;
      if (elect_one()) {
        const uint64_t da0 = desc_mn32(sx + (uint32_t)(r * p.W) * 128, 128), db0 = desc_mn32(sdy, kBK * 128);
#pragma unroll 4
        for (int k = 0; k < kBK / 8; ++k)       // 8 pixels (k-rows) per MMA = 1024 B further on = 64 descriptor units
          umma<true>(tacc, da0 + (uint64_t)(k * 64), db0 + (uint64_t)(k * 64), idesc, (i | k) != 0);
        umma_commit(&bars->empty[st]);
        if (i == n - 1) umma_commit(&bars->acc_full);
      }
      __syncwarp();
    }
  } else if (warp >= 4) {
    // ---- during the main loop: column sums of the dy tiles (bias gradient).  Thread = (16-byte unit u of the row,
    // row group rg); the TMA wrote the tile with the 32-byte-atom swizzle: physical unit = u ^ ((row & 3) << 1)
    const int t = (warp - 4) * 32 + lane;
    const int upr = p.Cout / 4, u = t % upr, rg = t / upr, nrg = 128 / upr;
    float4 bsum = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int i = 0; i < n; ++i) {
      const int st = i % nstages, ph = (i / nstages) & 1;
      mbar_wait(&bars->full[st], ph);
      const uint8_t* base = smem + st * stage_bytes + x_bytes + (u >> 3) * (kBK * 128);
#pragma unroll 4
      for (int row = rg; row < kBK; row += nrg) {
        const float4 v = *reinterpret_cast<const float4*>(base + row * 128 + (((u & 7) ^ ((row & 3) << 1)) << 4));
        bsum.x += v.x; bsum.y += v.y; bsum.z += v.z; bsum.w += v.w;
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&bars->empty[st]);
    }
    // ---- epilogue (once): accumulator r, lane = s*32 + c, column = k  ->  scratch[cta][k][(r*3+s)*32 + c]
    const int lg = warp & 3;                                   // lanes lg*32 .. +31  <=>  tap column s = lg, channel c = lane
    float* out = p.scratch + (size_t)blockIdx.x * p.Cout * 289;
    if (n > 0) { mbar_wait(&bars->acc_full, 0); tc_fence_after(); }
    // all MMAs have retired: the stages are free, use stage 0 to add the row groups of the bias sums
    float4* red = reinterpret_cast<float4*>(smem);
    red[rg * upr + u] = bsum;
    asm volatile("bar.sync 1, 128;" ::: "memory");
    if (t < p.Cout) {
      const float* rf = reinterpret_cast<const float*>(smem);
      float sum = 0.f;
      for (int g = 0; g < nrg; ++g) sum += rf[g * p.Cout + t];
      out[p.Cout * 288 + t] = sum;
    }
    if (lg < 3) {
      for (int r = 0; r < 3; ++r) {
        for (int c0 = 0; c0 < p.Cout; c0 += 32) {
          uint32_t v[32];
          if (n > 0) { tmem_ld32(tmem + ((uint32_t)(lg * 32) << 16) + r * 64 + c0, v); tmem_ld_wait(); }
          else {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = 0;
          }
          float* o = out + (r * 3 + lg) * 32 + lane;
#pragma unroll
          for (int j = 0; j < 32; ++j) o[(size_t)(c0 + j) * 288] = __uint_as_float(v[j]);    // lanes = consecutive c: 128-byte stores
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<kTmem>(tmem);
}

PFN_cuTensorMapEncodeTiled wg_encode() {
  static PFN_cuTensorMapEncodeTiled fn = nullptr;
  if (!fn) {
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void**)&fn, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) fn = nullptr;
  }
  return fn;
}
int wg_map(CUtensorMap* m, const void* base, long rows, long cols, int box_cols, int box_rows) {
  auto enc = wg_encode();
  if (!enc) { dcp_set_error("cuTensorMapEncodeTiled unavailable"); return -1; }
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)cols * 4};
  cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { dcp_set_error("flat wgrad: cuTensorMapEncodeTiled(rows=%ld cols=%ld box=%dx%d) failed: %d", rows, cols, box_cols, box_rows, (int)r); return -1; }
  return 0;
}
}  // namespace

extern "C" {

// x: [P, 32] fp32 flat pixels (pitch W); dy: [P, Cout] fp32 in the same frame; scratch: [num_ctas][Cout*288 + Cout] fp32 partial
// filters followed by partial bias gradients (finish with dcp_splitk_reduce2(scratch, dw, db, Cout*288, num_ctas, Cout*289)).
// 3x3 filters, Cin = 32.
// Exactly num_ctas partials are written (CTAs without pixels write zeros).
int dcp_conv_wgrad_flat_tc(const float* x, const float* dy, float* scratch, long P, int Cout, int W, int num_ctas, cudaStream_t s) {
  if (Cout != 32 && Cout != 64) { dcp_set_error("flat wgrad: Cout must be 32 or 64"); return -1; }
  WgradFlatParams p{};
  p.P = P; p.Cout = Cout; p.W = W; p.scratch = scratch;
  p.stages_total = (int)((P + kBK - 1) / kBK);
  const int halo_rows = (kBK + 2 * W + 3 + 7) / 8 * 8;
  if (halo_rows > 256) { dcp_set_error("flat wgrad: halo of %d rows exceeds one TMA box", halo_rows); return -1; }
  const int stage_bytes = halo_rows * 128 + (Cout / 32) * kBK * 128;
  int nstages = 3;
  const int budget = 200 * 1024;
  while (nstages > 1 && nstages * stage_bytes + (int)sizeof(WBars) + 1024 > budget) --nstages;
  const size_t smem = (size_t)nstages * stage_bytes + sizeof(WBars) + 1024;
  static bool attr = false;
  if (!attr) { DCP_CUDA_CHECK(cudaFuncSetAttribute(k_wgrad_flat, cudaFuncAttributeMaxDynamicSharedMemorySize, budget)); attr = true; }
  CUtensorMap mx, mdy;
  if (wg_map(&mx, x, P, 32, 32, halo_rows)) return -1;
  if (wg_map(&mdy, dy, P, Cout, 32, kBK)) return -1;
  if (num_ctas <= 0) { dcp_set_error("flat wgrad: num_ctas must be positive"); return -1; }
  DCP_CUDA_CHECK(dcp_launch_pdl(k_wgrad_flat, dim3(num_ctas), dim3(kThreads), smem, s, mx, mdy, p, nstages, halo_rows));
  return 0;
}

}  // extern "C"
