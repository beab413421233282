// Tensor-core GEMM and implicit-GEMM convolution for sm_100a: tcgen05.mma issued by one elected thread,
// operands staged in shared memory by TMA (128-byte swizzle), fp32 accumulator in TMEM, read back with
// tcgen05.ld for a fused bias / ReLU / cast epilogue.  One kernel serves
//   * Linear forward / data-grad / weight-grad (K8, K12 of the reference model, ref:main.py:39,43) through
//     the operand-major flags (K-major or MN-major per operand, so no transposes are materialised), and
//   * convolution forward and data-gradient (K3, ref:main.py:34) as an implicit GEMM: the A tile of filter
//     tap (r, s) is a 4-D TMA box {C, Wb, Hb, Nb} of the channels-last activation shifted by (r, s); TMA's
//     out-of-bounds zero fill implements the padding.
// CTA = 128 x BN output tile; warp 0 = TMA producer, warp 1 = MMA issuer (+ TMEM allocator), warps 2-5 =
// epilogue (each owns 32 TMEM lanes).  4-stage mbarrier pipeline, 128-byte-wide K slabs (32 tf32 or
// 64 bf16 elements).  gridDim.z splits K (fp32 atomics epilogue).
#include <cudaTypedefs.h>

#include <cstdlib>

#include "ops/tc_common.cuh"

struct TcGemmParams {
  int M, N;                 // logical output extent (rows, cols)
  int k_stages;             // number of 128-byte K slabs in total
  int BN;                   // N tile (32 / 64 / 128 <= kMaxBN)
  int b_bytes;              // bytes of one B stage (>= BN * 128; MN-major B is loaded in whole 128-byte chunks)
  int a_mn, b_mn;           // operand majors
  // conv = 1, implicit conv forward / data-grad (A via 4-D map): stage s -> tap = s / cchunks, channel chunk = s % cchunks
  // conv = 2, conv weight-grad: K runs over boxes of output pixels; A chunk u of accumulator group g is the
  //           activation box shifted by filter tap (unit / cchunks) for channel chunk (unit % cchunks); B is the dy box
  int conv, S, cchunks, Cin, pad, Wb, Hb, Nb, tiles_w, tiles_h, Ho, Wo, Nimg;
  int cstride;                  // convolution stride (1 or 2): the activation map samples every cstride-th pixel (TMA elementStrides)
  int stages;                   // pipeline depth (<= kStages)
  int x3;                       // fp32 only: 3xTF32 split (hi*hi + lo*hi + hi*lo) -> fp32-class accuracy
  int groups, units, a_bytes;   // accumulator groups per CTA (1 except wgrad), (tap, chunk) units in total, bytes of one A stage
  // epilogue
  void* out; long ldc; const void* bias; int relu, atomic, out_bf16, bias_bf16;
  long partial_stride;      // > 0: split z writes its partial tile to out + z * partial_stride with plain stores (reduced by dcp_splitk_reduce)
  long long* dbg;           // optional clock64 trace of CTA (0,0,0) (-DDCP_TRACE builds only)
};

namespace {
using namespace tc;

constexpr int kBM = 128, kStages = 4, kMaxBN = 128;
constexpr int kMaxGroups = 3;                 // wgrad: up to 3 accumulator groups (3 x 16 KB of A per stage)
constexpr int kABytes = kBM * 128;            // 16 KB per stage
constexpr int kThreads = 320;                // warp 0 TMA, warp 1 MMA, warps 2-9 operand split (3xTF32) + epilogue
constexpr int kEpiWarps = 8;

struct __align__(8) Pipe {
  uint64_t full[kStages], empty[kStages], split[kStages], acc_full;
  uint32_t tmem_base;
};

template <bool kTF32, int kTmemCols>
__global__ void __launch_bounds__(kThreads, 1)
k_tc_gemm(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b, const TcGemmParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // 1024-byte alignment is required by the 128-byte swizzle atoms
  uint8_t* smem = align_smem_1024(smem_raw);
  const int b_bytes = p.b_bytes;
  const int a_bytes = p.a_bytes;
  const int half_bytes = a_bytes + b_bytes;                 // [A][B]; in x3 mode followed by [A_lo][B_lo]
  const int stage_bytes = p.x3 ? 2 * half_bytes : half_bytes;
  const int nstages = p.stages;
  Pipe* pipe = reinterpret_cast<Pipe*>(smem + nstages * stage_bytes);
  const int warp = warp_idx_uniform(), lane = threadIdx.x & 31;
#ifdef DCP_TRACE      // clock64 trace of CTA (0,0,0) (build with DCP_NVCC_EXTRA=-DDCP_TRACE); compiled out of the production library
  long long* dbg = (p.dbg && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0) ? p.dbg : nullptr;
#else
  constexpr long long* dbg = nullptr;
#endif
  if (dbg && threadIdx.x == 0) dbg[0] = clock64();
  constexpr int kElem = kTF32 ? 4 : 2;
  constexpr int kBK = 128 / kElem;             // elements per K slab
  constexpr int kUmmaK = 32 / kElem;           // elements per MMA
  constexpr int kChunk = 128 / kElem;          // M/N elements per 128-byte chunk (MN-major)

  // K range of this split
  const int per = (p.k_stages + gridDim.z - 1) / gridDim.z;
  const int ks0 = blockIdx.z * per, ks1 = min(p.k_stages, ks0 + per);
  const int nst = ks1 - ks0;

  griddep_launch();
  if (warp == 0 && lane == 0) { tma_prefetch_desc(&map_a); tma_prefetch_desc(&map_b); }
  if (warp == 1) {
    if (lane == 0) {
      for (int i = 0; i < kStages; ++i) { mbar_init(&pipe->full[i], 1); mbar_init(&pipe->empty[i], 1); mbar_init(&pipe->split[i], kEpiWarps); }
      mbar_init(&pipe->acc_full, 1);
      mbar_fence_init();
    }
    __syncwarp();
    tmem_alloc<kTmemCols>(&pipe->tmem_base);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *((volatile uint32_t*)&pipe->tmem_base);
  griddep_wait();                      // prologue done; operands / outputs of earlier kernels are touched from here on
  if (dbg && threadIdx.x == 0) dbg[1] = clock64();

  const int m_tile = blockIdx.x, n_tile = blockIdx.y;
  // conv: decode the tile into (image group, box row, box col)
  int n0 = 0, h0 = 0, w0 = 0;
  if (p.conv == 1) {
    int t = m_tile;
    const int tw = t % p.tiles_w; t /= p.tiles_w;
    const int th = t % p.tiles_h; t /= p.tiles_h;
    n0 = t * p.Nb; h0 = th * p.Hb; w0 = tw * p.Wb;
  }

  if (nst > 0) {
    if (warp == 0) {
      // ================= TMA producer: the whole warp walks the loop and stays converged (it reaches the closing
      // __syncthreads as one), lane 0 issues =================
      for (int i = 0; i < nst; ++i) {
        const int st = i % nstages, ph = (i / nstages) & 1;
        mbar_wait(&pipe->empty[st], ph ^ 1);
        uint8_t* sa = smem + st * stage_bytes;
        uint8_t* sb = sa + a_bytes;
        const int ks = ks0 + i;
        if (lane != 0) { __syncwarp(); continue; }
        if (p.conv == 2) {
          // weight-grad: this stage = one box of kBK output pixels
          int t = ks;
          const int tw = t % p.tiles_w; t /= p.tiles_w;
          const int th = t % p.tiles_h; t /= p.tiles_h;
          const int pn = t * p.Nb, ph0 = th * p.Hb, pw0 = tw * p.Wb;
          constexpr int kUnitsPerGroup = kBM / kChunk;
          const int u0 = m_tile * p.groups * kUnitsPerGroup;
          const int nu = min(p.units - u0, p.groups * kUnitsPerGroup);
          const int nbc = (p.BN + kChunk - 1) / kChunk;
          mbar_expect_tx(&pipe->full[st], (nu + nbc) * (kBK * 128));
          for (int j = 0; j < nbc; ++j)
            tma_load_4d(sb + j * (kBK * 128), &map_b, &pipe->full[st], n_tile * p.BN + j * kChunk, pw0, ph0, pn);
          for (int u = 0; u < nu; ++u) {
            const int unit = u0 + u, tap = unit / p.cchunks, cc = unit % p.cchunks;
            tma_load_4d(sa + u * (kBK * 128), &map_a, &pipe->full[st], cc * kChunk, pw0 * p.cstride - p.pad + tap % p.S, ph0 * p.cstride - p.pad + tap / p.S, pn);
          }
          __syncwarp();
          continue;
        }
        mbar_expect_tx(&pipe->full[st], kABytes + b_bytes);
        int kb = ks * kBK;  // K coordinate for B (and for GEMM A)
        if (p.conv) {
          const int tap = ks / p.cchunks, cc = ks % p.cchunks;
          const int r = tap / p.S, s = tap % p.S;
          tma_load_4d(sa, &map_a, &pipe->full[st], cc * kBK, w0 * p.cstride - p.pad + s, h0 * p.cstride - p.pad + r, n0);
          kb = tap * p.Cin + cc * kBK;
        } else if (!p.a_mn) {
          tma_load_2d(sa, &map_a, &pipe->full[st], kb, m_tile * kBM);
        } else {
          for (int j = 0; j < kBM / kChunk; ++j)
            tma_load_2d(sa + j * (kBK * 128), &map_a, &pipe->full[st], m_tile * kBM + j * kChunk, kb);
        }
        if (!p.b_mn) {
          tma_load_2d(sb, &map_b, &pipe->full[st], kb, n_tile * p.BN);
        } else {
          for (int j = 0; j < (p.BN + kChunk - 1) / kChunk; ++j)
            tma_load_2d(sb + j * (kBK * 128), &map_b, &pipe->full[st], n_tile * p.BN + j * kChunk, kb);
        }
        if (dbg && i < 8) dbg[8 + i * 4 + 0] = clock64();
        __syncwarp();
      }
    } else if (warp == 1) {
      // ================= MMA issuer (uniform loop, one elected lane issues) =================
      const uint32_t idesc = make_idesc(kTF32, p.a_mn != 0, p.b_mn != 0, kBM, p.BN);
      for (int i = 0; i < nst; ++i) {
        const int st = i % nstages, ph = (i / nstages) & 1;
        mbar_wait(p.x3 ? &pipe->split[st] : &pipe->full[st], ph);
        tc_fence_after();
        if (dbg && i < 8 && lane == 0) dbg[8 + i * 4 + 1] = clock64();
        const uint32_t sa0 = smem_u32(smem + st * stage_bytes), sb = sa0 + a_bytes;
        if (elect_one()) {
        // descriptors: only the start-address field (16-byte units) moves between k-steps, so build one base per
        // operand and stage and add the step (MN-major: k-rows are 128 B apart; tf32 needs the BASE32B layout)
        constexpr uint32_t kMnSbo = kTF32 ? 512 : 1024, kMnLayout = kTF32 ? 1 : 2;
        const uint64_t db0 = p.b_mn ? make_smem_desc(sb, kBK * 128, kMnSbo, kMnLayout) : make_smem_desc(sb, 0, 1024);
        const uint64_t step_a = p.a_mn ? (uint64_t)((kUmmaK * 128) >> 4) : 2u, step_b = p.b_mn ? (uint64_t)((kUmmaK * 128) >> 4) : 2u;
        const uint64_t lo_step = (uint64_t)(half_bytes >> 4);
        for (int g = 0; g < p.groups; ++g) {
          const uint32_t sa = sa0 + g * kABytes;
          const uint32_t tmem_g = tmem + g * p.BN;
          const uint64_t da0 = p.a_mn ? make_smem_desc(sa, kBK * 128, kMnSbo, kMnLayout) : make_smem_desc(sa, 0, 1024);
#pragma unroll
          for (int j = 0; j < kBK / kUmmaK; ++j) {
            const uint64_t da = da0 + j * step_a, db = db0 + j * step_b;
            umma<kTF32>(tmem_g, da, db, idesc, (i | j) != 0);
            if (kTF32 && p.x3) {   // + A_lo * B_hi + A_hi * B_lo (the lo tiles sit half_bytes further on)
              umma<kTF32>(tmem_g, da + lo_step, db, idesc, 1);
              umma<kTF32>(tmem_g, da, db + lo_step, idesc, 1);
            }
          }
        }
        umma_commit(&pipe->empty[st]);   // smem slot reusable once these MMAs retire
        if (i == nst - 1) umma_commit(&pipe->acc_full);      // accumulator complete
        }
        __syncwarp();
        if (dbg && i < 8 && lane == 0) dbg[8 + i * 4 + 2] = clock64();
      }
    }
  }
  if (warp >= 2) {
    // ================= epilogue: TMEM -> registers -> global =================
    const int lg = warp & 3;                       // TMEM lane group this warp may access
    const int row = lg * 32 + lane;                // row inside the tile
    if (kTF32 && p.x3) {
      // 3xTF32: while the pipeline runs, these 256 threads split every landed stage in place into
      // hi = x with the low 13 mantissa bits cleared (exact in tf32) and lo = x - hi (exact in fp32).
      // The split is element-wise, so it is independent of the swizzle the TMA applied.
      const int t = (warp - 2) * 32 + lane;
      for (int i = 0; i < nst; ++i) {
        const int st = i % nstages, ph = (i / nstages) & 1;
        mbar_wait(&pipe->full[st], ph);
        if (dbg && i < 8 && t == 0) dbg[8 + i * 4 + 3] = clock64();
        uint4* hi = reinterpret_cast<uint4*>(smem + st * stage_bytes);
        uint4* lo = reinterpret_cast<uint4*>(smem + st * stage_bytes + half_bytes);
#pragma unroll 4
        for (int v = t; v < half_bytes / 16; v += kEpiWarps * 32) {
          uint4 x = hi[v], h, l;
          h.x = x.x & 0xFFFFE000u; h.y = x.y & 0xFFFFE000u; h.z = x.z & 0xFFFFE000u; h.w = x.w & 0xFFFFE000u;
          l.x = __float_as_uint(__uint_as_float(x.x) - __uint_as_float(h.x)); l.y = __float_as_uint(__uint_as_float(x.y) - __uint_as_float(h.y));
          l.z = __float_as_uint(__uint_as_float(x.z) - __uint_as_float(h.z)); l.w = __float_as_uint(__uint_as_float(x.w) - __uint_as_float(h.w));
          hi[v] = h; lo[v] = l;
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy writes -> visible to the tensor core
        __syncwarp();
        if (lane == 0) mbar_arrive(&pipe->split[st]);
      }
    }
    if (nst > 0) {
      mbar_wait(&pipe->acc_full, 0);
      tc_fence_after();
    }
    if (dbg && warp == 2 && lane == 0) dbg[3] = clock64();
    if (p.conv == 2 && warp < 6) {
      // weight-grad: row = (unit-in-group, channel-in-chunk), column = output channel; dw[k][tap][c] += acc
      constexpr int kUnitsPerGroup = kBM / kChunk;
      for (int g = 0; g < p.groups; ++g) {
        const int unit = (m_tile * p.groups + g) * kUnitsPerGroup + row / kChunk;
        const bool valid = unit < p.units;
        const int tap = unit / p.cchunks, cc = unit % p.cchunks;
        float* orow = (float*)p.out + (p.partial_stride ? (long)blockIdx.z * p.partial_stride : 0) + (long)tap * p.Cin + cc * kChunk + (row % kChunk);
        for (int c0 = 0; c0 < p.BN; c0 += 32) {
          uint32_t r[32];
          if (nst > 0) { tmem_ld32(tmem + ((uint32_t)(lg * 32) << 16) + g * p.BN + c0, r); tmem_ld_wait(); }
          if (!valid) continue;
          if (p.partial_stride) {
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              const int k = n_tile * p.BN + c0 + j;
              if (k < p.N) orow[(long)k * p.ldc] = nst > 0 ? __uint_as_float(r[j]) : 0.f;   // lanes = consecutive c: coalesced 128-B stores
            }
            continue;
          }
          if (nst <= 0) continue;
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const int k = n_tile * p.BN + c0 + j;
            if (k < p.N) atomicAdd(orow + (long)k * p.ldc, __uint_as_float(r[j]));   // lanes = consecutive c: coalesced
          }
        }
      }
    } else if (p.conv != 2) {
    // Every MMA has retired (acc_full), so the operand stages are dead: reuse them as the transpose buffer that turns
    // "one thread = one output row" (what tcgen05.ld delivers) into coalesced stores - 8 lanes x 16 B cover the 32
    // columns of a row, 4 rows per store instruction.  Warps 2-5 take the first half of the BN columns, 6-9 the second.
    const int ew = warp - 2;
    float* stg = reinterpret_cast<float*>(smem) + ew * (32 * 36 + 64);
    long* srow = reinterpret_cast<long*>(stg + 32 * 36);
    long out_row = -1;
    if (p.conv) {
      const int nl = row / (p.Hb * p.Wb), hl = (row / p.Wb) % p.Hb, wl = row % p.Wb;
      const int n = n0 + nl, h = h0 + hl, w = w0 + wl;
      if (n < p.Nimg && h < p.Ho && w < p.Wo) out_row = ((long)n * p.Ho + h) * p.Wo + w;
    } else {
      const long gm = (long)m_tile * kBM + row;
      if (gm < p.M) out_row = gm;
    }
    srow[lane] = out_row;
    const bool add_bias = p.bias != nullptr && blockIdx.z == 0 && p.partial_stride == 0;
    const long zoff = p.partial_stride ? (long)blockIdx.z * p.partial_stride : 0;
    const int cpw = p.BN >= 64 ? p.BN / 2 : 32;             // columns per warp set
    const int cbeg = (ew >> 2) * cpw, cend = min(p.BN, cbeg + cpw);
    for (int c0 = cbeg; c0 < cend; c0 += 32) {
      uint32_t r[32];
      if (nst > 0) {
        tmem_ld32(tmem + ((uint32_t)(lg * 32) << 16) + c0, r);
        tmem_ld_wait();
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) r[j] = 0;
      }
      const int gn0 = n_tile * p.BN + c0;
      const int ncol = min(32, p.N - gn0);
      if (ncol <= 0) break;
      if (add_bias || p.relu) {
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          const int gn = gn0 + j;
          float v = __uint_as_float(r[j]);
          if (add_bias && gn < p.N) v += p.bias_bf16 ? __bfloat162float(((const __nv_bfloat16*)p.bias)[gn]) : ((const float*)p.bias)[gn];
          if (p.relu) v = fmaxf(v, 0.f);
          r[j] = __float_as_uint(v);
        }
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) *reinterpret_cast<uint4*>(stg + lane * 36 + j * 4) = make_uint4(r[4 * j], r[4 * j + 1], r[4 * j + 2], r[4 * j + 3]);
      __syncwarp();
      const int c4 = (lane & 7) * 4, nv = min(4, ncol - c4);
#pragma unroll
      for (int it = 0; it < 8; ++it) {
        const int rl = it * 4 + (lane >> 3);
        const long orow = srow[rl];
        if (orow < 0 || nv <= 0) continue;
        const float4 v = *reinterpret_cast<const float4*>(stg + rl * 36 + c4);
        const float ve[4] = {v.x, v.y, v.z, v.w};
        if (p.atomic) {
          float* o = (float*)p.out + orow * p.ldc + gn0 + c4;
          for (int e = 0; e < nv; ++e) atomicAdd(o + e, ve[e]);
        } else if (p.out_bf16) {
          __nv_bfloat16* o = (__nv_bfloat16*)p.out + orow * p.ldc + gn0 + c4;
          if (nv == 4 && (reinterpret_cast<uintptr_t>(o) & 7) == 0) {
            const __nv_bfloat162 t0 = __floats2bfloat162_rn(v.x, v.y), t1 = __floats2bfloat162_rn(v.z, v.w);
            *reinterpret_cast<uint2*>(o) = make_uint2(*reinterpret_cast<const uint32_t*>(&t0), *reinterpret_cast<const uint32_t*>(&t1));
          } else {
            for (int e = 0; e < nv; ++e) o[e] = __float2bfloat16_rn(ve[e]);
          }
        } else {
          float* o = (float*)p.out + zoff + orow * p.ldc + gn0 + c4;
          if (nv == 4 && (reinterpret_cast<uintptr_t>(o) & 15) == 0) *reinterpret_cast<float4*>(o) = v;
          else for (int e = 0; e < nv; ++e) o[e] = ve[e];
        }
      }
      __syncwarp();
    }
    }
  }
  if (dbg && warp == 2 && lane == 0) dbg[4] = clock64();
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<kTmemCols>(tmem);
  if (dbg && threadIdx.x == 0) dbg[2] = clock64();
}

// ---- host: tensor maps ------------------------------------------------------------------------------
PFN_cuTensorMapEncodeTiled get_encode() {
  static PFN_cuTensorMapEncodeTiled fn = nullptr;
  if (!fn) {
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void**)&fn, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) fn = nullptr;
  }
  return fn;
}

// 2-D map over a row-major [rows][cols] matrix; box = {box_cols (<= 128 B), box_rows}
int make_map_2d(CUtensorMap* m, const void* base, int elem_bytes, long rows, long cols, long ld_elems, int box_cols, int box_rows,
                bool mn_major = false) {
  auto enc = get_encode();
  if (!enc) { dcp_set_error("cuTensorMapEncodeTiled unavailable"); return -1; }
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld_elems * elem_bytes};
  cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUtensorMapDataType dt = elem_bytes == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  // fp32 operands read MN-major must land in the 32-byte-atom swizzle (see tc_common.cuh)
  CUtensorMapSwizzle sw = (mn_major && elem_bytes == 4) ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B;
  CUresult r = enc(m, dt, 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { dcp_set_error("cuTensorMapEncodeTiled(2d rows=%ld cols=%ld ld=%ld box=%dx%d) failed: %d", rows, cols, ld_elems, box_cols, box_rows, (int)r); return -1; }
  return 0;
}
// 4-D map over channels-last [N][H][W][C]; box = {Cb, Wb, Hb, Nb}
// estride > 1: the box samples every estride-th pixel along W and H (a strided convolution's input window); the box still
// delivers Cb x Wb x Hb x Nb elements to shared memory
int make_map_4d(CUtensorMap* m, const void* base, int elem_bytes, int N, int H, int W, int C, int Cb, int Wb, int Hb, int Nb, bool mn_major = false,
                int estride = 1) {
  auto enc = get_encode();
  if (!enc) { dcp_set_error("cuTensorMapEncodeTiled unavailable"); return -1; }
  cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
  cuuint64_t strides[3] = {(cuuint64_t)C * elem_bytes, (cuuint64_t)W * C * elem_bytes, (cuuint64_t)H * W * C * elem_bytes};
  cuuint32_t box[4] = {(cuuint32_t)Cb, (cuuint32_t)(Wb * estride), (cuuint32_t)(Hb * estride), (cuuint32_t)Nb};
  cuuint32_t estr[4] = {1, (cuuint32_t)estride, (cuuint32_t)estride, 1};
  CUtensorMapDataType dt = elem_bytes == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  CUtensorMapSwizzle sw = (mn_major && elem_bytes == 4) ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B;
  CUresult r = enc(m, dt, 4, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { dcp_set_error("cuTensorMapEncodeTiled(4d N=%d H=%d W=%d C=%d box=%d,%d,%d,%d) failed: %d", N, H, W, C, Cb, Wb, Hb, Nb, (int)r); return -1; }
  return 0;
}

size_t smem_bytes(int stages, int a_bytes, int b_bytes, int x3 = 0) { return (size_t)stages * (a_bytes + b_bytes) * (x3 ? 2 : 1) + sizeof(Pipe) + 1024; }
constexpr int kMaxSmem = 200 * 1024;    // >= 8 x 4864 B of epilogue staging, which reuses the (dead) operand stages
int b_stage_bytes(int BN, int b_mn, int es) { int chunk = 128 / es, bk = 128 / es; return b_mn ? ((BN + chunk - 1) / chunk) * bk * 128 : BN * 128; }

template <bool kTF32, int kTmemCols> int launch(const CUtensorMap& ma, const CUtensorMap& mb, const TcGemmParams& p, dim3 grid, cudaStream_t s) {
  static bool attr_set = false;
  if (!attr_set) {
    DCP_CUDA_CHECK(cudaFuncSetAttribute(k_tc_gemm<kTF32, kTmemCols>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        kMaxSmem));
    attr_set = true;
  }
  TcGemmParams q = p;
#ifdef DCP_TRACE
  { const char* e = getenv("DCP_GEMM_DBG_PTR"); q.dbg = e ? (long long*)strtoull(e, nullptr, 0) : nullptr; }
#else
  q.dbg = nullptr;
#endif
  q.stages = kStages;
  while (q.stages > 2 && smem_bytes(q.stages, q.a_bytes, q.b_bytes, q.x3) > (size_t)kMaxSmem) --q.stages;
  if (smem_bytes(q.stages, q.a_bytes, q.b_bytes, q.x3) > (size_t)kMaxSmem) { dcp_set_error("tc kernel: stage of %d bytes does not fit shared memory", q.a_bytes + q.b_bytes); return -1; }
  DCP_CUDA_CHECK(dcp_launch_pdl(k_tc_gemm<kTF32, kTmemCols>, grid, dim3(kThreads), smem_bytes(q.stages, q.a_bytes, q.b_bytes, q.x3), s, ma, mb, q));
  return 0;
}

int pick_bn(int N) { return N <= 32 ? 32 : (N <= 64 ? 64 : 128); }

}  // namespace

// -1: DCP_GEMM_NARROW decides; 0 / 1 / 2: the caller decides for the launches that follow (two such GEMMs issued on two streams
// to run side by side want the wide tiles: 72 CTAs each, both resident at once) - see dcp_gemm_set_narrow
static int g_gemm_narrow_override = -1;

extern "C" {
// Host-side launch choice for wide, shallow fp32 / tf32 problems: 64-column tiles (twice the CTAs) when 1, 128-column when 0,
// the DCP_GEMM_NARROW default when -1.  Affects the launches issued after the call from this process.
void dcp_gemm_set_narrow(int v) { g_gemm_narrow_override = v; }


// C[M,N] (+)= op(A) . op(B)^T with fp32 accumulation on tcgen05.
//   a_mn = 0: A is [M][K] row-major (lda);   a_mn = 1: A is [K][M] row-major (lda)
//   b_mn = 0: B is [N][K] row-major (ldb);   b_mn = 1: B is [K][N] row-major (ldb)
// dtype: DCP_F32 (tf32 multiply) or DCP_BF16.  out: fp32 or bf16 (out_bf16), ldc elements.  ksplit > 1 => `out`
// must be fp32 and zero-initialised (atomic accumulation; relu is rejected).  K must be a multiple of the 128-byte slab.
// x3 (fp32 only): 3xTF32 - every operand is split into a tf32-exact high part and a low remainder and the kernel
// accumulates hi*hi + lo*hi + hi*lo, which restores fp32-class accuracy (what cuBLAS gives the reference's Linear
// layers with allow_tf32=False) at three times the MMA count.
int dcp_tc_gemm_ex(const void* A, const void* B, void* out, const void* bias, int M, int N, int K, long lda, long ldb, long ldc, int a_mn, int b_mn,
                   int dtype, int out_bf16, int bias_bf16, int relu, int ksplit, int x3, cudaStream_t s);
int dcp_gemm_bf16_persistent(const void* A, const void* B, void* out, const void* bias, int M, int N, int K, long lda, long ldb, long ldc,
                             int a_mn, int b_mn, int out_bf16, int bias_bf16, int relu, int ksplit, int num_ctas, cudaStream_t s);
static bool bf16_persistent_enabled() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("DCP_GEMM_PERSISTENT"); v = (e && e[0] == '0') ? 0 : 1; }
  return v != 0;
}
int dcp_tc_gemm(const void* A, const void* B, void* out, const void* bias, int M, int N, int K, long lda, long ldb, long ldc, int a_mn, int b_mn,
                int dtype, int out_bf16, int bias_bf16, int relu, int ksplit, cudaStream_t s) {
  return dcp_tc_gemm_ex(A, B, out, bias, M, N, K, lda, ldb, ldc, a_mn, b_mn, dtype, out_bf16, bias_bf16, relu, ksplit, 0, s);
}
int dcp_tc_gemm_ex(const void* A, const void* B, void* out, const void* bias, int M, int N, int K, long lda, long ldb, long ldc, int a_mn, int b_mn,
                   int dtype, int out_bf16, int bias_bf16, int relu, int ksplit, int x3, cudaStream_t s) {
  const int es = dtype == DCP_F32 ? 4 : 2, bk = 128 / es;
  if (dtype != DCP_F32 && dtype != DCP_BF16) { dcp_set_error("tc gemm: dtype must be f32 or bf16"); return -1; }
  if (K % bk) { dcp_set_error("tc gemm: K=%d must be a multiple of %d", K, bk); return -1; }
  // ksplit < -1: split-K through a scratch buffer: `out` is then a [|ksplit|][M*ldc] fp32 scratch that receives the partial
  // tiles with plain stores (no atomics); finish with dcp_splitk_reduce (which also adds the bias)
  const bool partial = ksplit < -1;
  if (partial) ksplit = -ksplit;
  // bf16: the persistent 128 x 256 kernel (gemm_bf16_persistent.cu) unless DCP_GEMM_PERSISTENT=0 or atomics are asked for
  if (dtype == DCP_BF16 && (ksplit <= 1 || partial) && bf16_persistent_enabled())
    return dcp_gemm_bf16_persistent(A, B, out, bias, M, N, K, lda, ldb, ldc, a_mn, b_mn, out_bf16, bias_bf16, relu, ksplit < 1 ? 1 : ksplit, 0, s);
  if (ksplit > 1 && (relu || out_bf16)) { dcp_set_error("tc gemm: split-K needs an fp32 output and no ReLU"); return -1; }
  TcGemmParams p{};
  p.M = M; p.N = N; p.k_stages = K / bk; p.BN = pick_bn(N); p.a_mn = a_mn; p.b_mn = b_mn;
  {
    // wide, shallow problems (fc1's weight / data gradient: 128 x 9216, K = 128) are pure streaming: with 128-column tiles only
    // 72 CTAs exist and half the SMs (and their share of the memory system) idle; 64-column tiles double the CTAs
    static const int narrow_env = [] { const char* e = getenv("DCP_GEMM_NARROW"); return e ? atoi(e) : 1; }();
    const int narrow = g_gemm_narrow_override >= 0 ? g_gemm_narrow_override : narrow_env;
    const long tiles = (long)((M + kBM - 1) / kBM) * ((N + p.BN - 1) / p.BN) * (ksplit < 1 ? 1 : ksplit);
    if (narrow && p.BN == 128 && tiles < 100) p.BN = narrow >= 2 ? 32 : 64;
  }
  p.b_bytes = b_stage_bytes(p.BN, b_mn, es); p.a_bytes = kABytes; p.groups = 1; p.x3 = (x3 && dtype == DCP_F32) ? 1 : 0;
  p.out = out; p.ldc = ldc; p.bias = bias; p.relu = relu; p.atomic = ksplit > 1 && !partial; p.out_bf16 = out_bf16; p.bias_bf16 = bias_bf16;
  p.partial_stride = partial ? (long)M * ldc : 0;
  CUtensorMap ma, mb;
  if (!a_mn) { if (make_map_2d(&ma, A, es, M, K, lda, bk, kBM)) return -1; }
  else { if (make_map_2d(&ma, A, es, K, M, lda, bk, bk, true)) return -1; }
  if (!b_mn) { if (make_map_2d(&mb, B, es, N, K, ldb, bk, p.BN)) return -1; }
  else { if (make_map_2d(&mb, B, es, K, N, ldb, bk, bk, true)) return -1; }
  dim3 grid((M + kBM - 1) / kBM, (N + p.BN - 1) / p.BN, ksplit < 1 ? 1 : ksplit);
  return dtype == DCP_F32 ? launch<true, 128>(ma, mb, p, grid, s) : launch<false, 128>(ma, mb, p, grid, s);
}

// Implicit-GEMM convolution (stride 1): y[n,ho,wo,k] = sum_{r,s,c} x[n, ho-pad+r, wo-pad+s, c] * w[k][(r*S+s)*C + c]
//   x: channels-last [N,H,W,C]; wmat: [Kout][R*S*C] row-major; y: channels-last [N,Ho,Wo,Kout] (fp32 or bf16).
// C must be a multiple of the 128-byte slab (32 f32 / 64 bf16).  (Wb,Hb,Nb): box of output pixels per CTA, product 128.
int dcp_tc_conv_strided(const void* x, const void* wmat, void* y, const void* bias, int N, int H, int W, int C, int Kout, int R, int S, int pad,
                        int stride, int Ho, int Wo, int Wb, int Hb, int Nb, int dtype, int out_bf16, int bias_bf16, int relu, cudaStream_t s);
int dcp_tc_conv(const void* x, const void* wmat, void* y, const void* bias, int N, int H, int W, int C, int Kout, int R, int S, int pad, int Ho,
                int Wo, int Wb, int Hb, int Nb, int dtype, int out_bf16, int bias_bf16, int relu, cudaStream_t s) {
  return dcp_tc_conv_strided(x, wmat, y, bias, N, H, W, C, Kout, R, S, pad, 1, Ho, Wo, Wb, Hb, Nb, dtype, out_bf16, bias_bf16, relu, s);
}
// Same with a convolution stride of 1 or 2: y[n,ho,wo,k] = sum x[n, ho*stride-pad+r, wo*stride-pad+s, c] * w[k][(r*S+s)*C + c].
// The strided input window is one TMA box with elementStrides = stride along W and H.
int dcp_tc_conv_strided(const void* x, const void* wmat, void* y, const void* bias, int N, int H, int W, int C, int Kout, int R, int S, int pad,
                        int stride, int Ho, int Wo, int Wb, int Hb, int Nb, int dtype, int out_bf16, int bias_bf16, int relu, cudaStream_t s) {
  const int es = dtype == DCP_F32 ? 4 : 2, bk = 128 / es;
  if (stride < 1 || stride > 2 || Wb * stride > 256 || Hb * stride > 256) { dcp_set_error("tc conv: stride %d / box %dx%d not supported", stride, Wb, Hb); return -1; }
  if (C % bk) { dcp_set_error("tc conv: C=%d must be a multiple of %d", C, bk); return -1; }
  if (Wb * Hb * Nb != kBM) { dcp_set_error("tc conv: box %dx%dx%d must cover %d pixels", Wb, Hb, Nb, kBM); return -1; }
  TcGemmParams p{};
  p.N = Kout; p.BN = pick_bn(Kout); p.b_bytes = p.BN * 128; p.a_bytes = kABytes; p.groups = 1; p.conv = 1; p.S = S; p.cchunks = C / bk; p.Cin = C; p.pad = pad;
  p.k_stages = R * S * p.cchunks; p.cstride = stride;
  p.Wb = Wb; p.Hb = Hb; p.Nb = Nb; p.tiles_w = (Wo + Wb - 1) / Wb; p.tiles_h = (Ho + Hb - 1) / Hb; p.Ho = Ho; p.Wo = Wo; p.Nimg = N;
  p.out = y; p.ldc = Kout; p.bias = bias; p.relu = relu; p.out_bf16 = out_bf16; p.bias_bf16 = bias_bf16;
  p.M = p.tiles_w * p.tiles_h * ((N + Nb - 1) / Nb) * kBM;
  CUtensorMap ma, mb;
  if (make_map_4d(&ma, x, es, N, H, W, C, bk, Wb, Hb, Nb, false, stride)) return -1;
  if (make_map_2d(&mb, wmat, es, Kout, (long)R * S * C, (long)R * S * C, bk, p.BN)) return -1;
  dim3 grid(p.M / kBM, (Kout + p.BN - 1) / p.BN, 1);
  return dtype == DCP_F32 ? launch<true, 128>(ma, mb, p, grid, s) : launch<false, 128>(ma, mb, p, grid, s);
}

// Convolution weight gradient (stride 1) on tcgen05: dw[k][r][s][c] += sum_pixels dy[n,ho,wo,k] * x[n, ho-pad+r, wo-pad+s, c].
// Both operands are read MN-major straight from the channels-last tensors (no im2col, no transposes): the K
// dimension of the MMA runs over boxes of (Wb x Hb x Nb) = 32 (f32) / 64 (bf16) output pixels, the A operand of
// accumulator group g is made of 128-byte channel chunks taken from the activation shifted by different filter
// taps, B is the matching box of dy.  gridDim.z splits the pixel boxes; partial sums meet in fp32 atomics, so
// dw32 ([Kout][R*S*C] fp32) is zeroed here.
int dcp_tc_conv_wgrad_strided(const void* dy, const void* x, float* dw32, int N, int H, int W, int C, int Kout, int R, int S, int pad, int stride,
                              int Ho, int Wo, int Wb, int Hb, int Nb, int dtype, int zsplit, cudaStream_t s);
int dcp_tc_conv_wgrad(const void* dy, const void* x, float* dw32, int N, int H, int W, int C, int Kout, int R, int S, int pad, int Ho, int Wo,
                      int Wb, int Hb, int Nb, int dtype, int zsplit, cudaStream_t s) {
  return dcp_tc_conv_wgrad_strided(dy, x, dw32, N, H, W, C, Kout, R, S, pad, 1, Ho, Wo, Wb, Hb, Nb, dtype, zsplit, s);
}
// Same for a convolution stride of 1 or 2 (x is read at ho*stride-pad+r, wo*stride-pad+s).
int dcp_tc_conv_wgrad_strided(const void* dy, const void* x, float* dw32, int N, int H, int W, int C, int Kout, int R, int S, int pad, int stride,
                              int Ho, int Wo, int Wb, int Hb, int Nb, int dtype, int zsplit, cudaStream_t s) {
  const int es = dtype == DCP_F32 ? 4 : 2, chunk = 128 / es;
  if (stride < 1 || stride > 2) { dcp_set_error("tc wgrad: stride %d not supported", stride); return -1; }
  if (C % chunk) { dcp_set_error("tc wgrad: C=%d must be a multiple of %d", C, chunk); return -1; }
  if (Wb * Hb * Nb != chunk) { dcp_set_error("tc wgrad: pixel box must hold %d pixels", chunk); return -1; }
  TcGemmParams p{};
  p.conv = 2; p.cstride = stride; p.N = Kout; p.BN = pick_bn(Kout); p.a_mn = 1; p.b_mn = 1; p.S = S; p.cchunks = C / chunk; p.Cin = C; p.pad = pad;
  p.units = R * S * p.cchunks;
  const int upg = kBM / chunk;                       // units per accumulator group
  const int groups_total = (p.units + upg - 1) / upg;
  p.groups = groups_total < kMaxGroups ? groups_total : kMaxGroups;
  if (p.groups * p.BN > 512) p.groups = 512 / p.BN;
  p.a_bytes = p.groups * kABytes;
  p.b_bytes = b_stage_bytes(p.BN, 1, es);
  p.Wb = Wb; p.Hb = Hb; p.Nb = Nb; p.tiles_w = (Wo + Wb - 1) / Wb; p.tiles_h = (Ho + Hb - 1) / Hb; p.Ho = Ho; p.Wo = Wo; p.Nimg = N;
  p.k_stages = p.tiles_w * p.tiles_h * ((N + Nb - 1) / Nb);
  // zsplit < -1: partial mode - dw32 is a [|zsplit|][Kout*R*S*C] scratch, finish with dcp_splitk_reduce
  const bool partial = zsplit < -1;
  if (partial) zsplit = -zsplit;
  p.out = dw32; p.ldc = (long)R * S * C; p.atomic = 1;
  p.partial_stride = partial ? (long)Kout * R * S * C : 0;
  if (!partial) DCP_CUDA_CHECK(cudaMemsetAsync(dw32, 0, sizeof(float) * (size_t)Kout * R * S * C, s));
  CUtensorMap ma, mb;
  if (make_map_4d(&ma, x, es, N, H, W, C, chunk, Wb, Hb, Nb, true, stride)) return -1;
  if (make_map_4d(&mb, dy, es, N, Ho, Wo, Kout, chunk, Wb, Hb, Nb, true)) return -1;
  const int gtiles = (groups_total + p.groups - 1) / p.groups;
  if (zsplit < 1) zsplit = 1;
  if (zsplit > p.k_stages) zsplit = p.k_stages;
  dim3 grid(gtiles, (Kout + p.BN - 1) / p.BN, zsplit);
  return dtype == DCP_F32 ? launch<true, 512>(ma, mb, p, grid, s) : launch<false, 512>(ma, mb, p, grid, s);
}

}  // extern "C"

// Weight repack for the data-gradient: wt[c][(r', s', k)] = w[k][R-1-r'][S-1-s'][c]  (w is [K][R][S][C]).
namespace {
template <typename T> __global__ void k_flip_weights(const T* __restrict__ w, T* __restrict__ wt, int K, int R, int S, int C) {
  griddep_launch();
  griddep_wait();
  const long total = (long)K * R * S * C;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    // i indexes wt: ((c * R + r') * S + s') * K + k
    int k = i % K; long t = i / K; int s2 = t % S; t /= S; int r2 = t % R; int c = t / R;
    wt[i] = w[(((long)k * R + (R - 1 - r2)) * S + (S - 1 - s2)) * C + c];
  }
}
}  // namespace
// out[i] = (bias ? bias[i % ncols] : 0) + sum_z scratch[z][i]   - finishes a split-K done through scratch
namespace {
// CTA = 32 float4 columns x 8 split lanes: the (long) chain of dependent loads over the splits is cut 8 ways and the
// grid stays wide even for a 128 x 128 output
__global__ void __launch_bounds__(256) k_splitk_reduce(const float* __restrict__ scratch, float* __restrict__ out, const float* __restrict__ bias,
                                                       int nsplit, long n, int ncols, float* __restrict__ out2, long n1) {
  __shared__ float4 red[8][33];
  griddep_launch();
  griddep_wait();
  const int vx = threadIdx.x & 31, zl = threadIdx.x >> 5;
  const long v = (long)blockIdx.x * 32 + vx, n4 = n / 4;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  if (v < n4) {
#pragma unroll 4
    for (int z = zl; z < nsplit; z += 8) {
      const float4 x = *reinterpret_cast<const float4*>(scratch + (long)z * n + v * 4);
      acc.x += x.x; acc.y += x.y; acc.z += x.z; acc.w += x.w;
    }
  }
  red[zl][vx] = acc;
  __syncthreads();
  if (zl == 0 && v < n4) {
#pragma unroll
    for (int i = 1; i < 8; ++i) { const float4 x = red[i][vx]; acc.x += x.x; acc.y += x.y; acc.z += x.z; acc.w += x.w; }
    if (bias) { const int c = (int)((v * 4) % ncols); acc.x += bias[c]; acc.y += bias[c + 1]; acc.z += bias[c + 2]; acc.w += bias[c + 3]; }
    *reinterpret_cast<float4*>(v * 4 < n1 ? out + v * 4 : out2 + (v * 4 - n1)) = acc;
  }
}
// few splits: one thread per output vector, the splits are consecutive independent loads (fully coalesced, no shared memory)
__global__ void __launch_bounds__(256) k_splitk_reduce_few(const float* __restrict__ scratch, float* __restrict__ out, const float* __restrict__ bias,
                                                           int nsplit, long n, int ncols, float* __restrict__ out2, long n1) {
  griddep_launch();
  griddep_wait();
  const long n4 = n / 4;
  for (long v = (long)blockIdx.x * blockDim.x + threadIdx.x; v < n4; v += (long)gridDim.x * blockDim.x) {
    float4 acc = *reinterpret_cast<const float4*>(scratch + v * 4);
    for (int z = 1; z < nsplit; ++z) {
      const float4 x = *reinterpret_cast<const float4*>(scratch + (long)z * n + v * 4);
      acc.x += x.x; acc.y += x.y; acc.z += x.z; acc.w += x.w;
    }
    if (bias) { const int c = (int)((v * 4) % ncols); acc.x += bias[c]; acc.y += bias[c + 1]; acc.z += bias[c + 2]; acc.w += bias[c + 3]; }
    *reinterpret_cast<float4*>(v * 4 < n1 ? out + v * 4 : out2 + (v * 4 - n1)) = acc;
  }
}
}  // namespace
// n and ncols must be multiples of 4.
extern "C" int dcp_splitk_reduce(const float* scratch, float* out, const float* bias, int nsplit, long n, int ncols, cudaStream_t s) {
  if ((n & 3) || (ncols & 3)) { dcp_set_error("splitk reduce: n and ncols must be multiples of 4"); return -1; }
  if (nsplit <= 4) {
    long g = (n / 4 + 255) / 256; if (g > 148 * 16) g = 148 * 16; if (g < 1) g = 1;
    DCP_CUDA_CHECK(dcp_launch_pdl(k_splitk_reduce_few, dim3((unsigned)g), dim3(256), 0, s, scratch, out, bias, nsplit, n, ncols, (float*)nullptr, n));
    return 0;
  }
  DCP_CUDA_CHECK(dcp_launch_pdl(k_splitk_reduce, dim3((unsigned)((n / 4 + 31) / 32)), dim3(256), 0, s, scratch, out, bias, nsplit, n, ncols, (float*)nullptr, n));
  return 0;
}
// Same reduction with the summed vector going to two places: elements [0, n1) to `out`, [n1, n) to `out2` (n1 % 4 == 0).
extern "C" int dcp_splitk_reduce2(const float* scratch, float* out, float* out2, long n1, int nsplit, long n, cudaStream_t s) {
  if ((n & 3) || (n1 & 3) || n1 > n) { dcp_set_error("splitk reduce2: n and n1 must be multiples of 4, n1 <= n"); return -1; }
  DCP_CUDA_CHECK(dcp_launch_pdl(k_splitk_reduce, dim3((unsigned)((n / 4 + 31) / 32)), dim3(256), 0, s, scratch, out, (const float*)nullptr, nsplit, n, 4, out2, n1));
  return 0;
}

extern "C" int dcp_flip_weights(const void* w, void* wt, int K, int R, int S, int C, int dtype, cudaStream_t s) {
  long total = (long)K * R * S * C;
  DISPATCH_ACT(dtype, DCP_CUDA_CHECK(dcp_launch_pdl(k_flip_weights<T>, dim3(grid_1d(total, 256)), dim3(256), 0, s, (const T*)w, (T*)wt, K, R, S, C)));
  return 0;
}
