// Fused reduce_scatter -> optimizer update -> all_gather in ONE kernel over NVLink peer memory
// (north-star C11; a ZeRO-1-shaped step - not in the reference, which calls optimizer.step() on replicated
// state after DDP's all-reduce, ref:main.py:62-63).
//
// Layout: gradients sit in a flat bucket and parameters in a flat buffer with the same element order, both inside
// the symmetric heap at the same offsets on every rank.  Rank r owns elements [r*n/W, (r+1)*n/W) of the bucket and
// keeps optimizer state only for that shard.  Per 16-byte vector of its shard a thread
//   1. pulls the W gradient copies straight from the peers' buckets (NVLink loads, all in flight together) and
//      averages them in registers                                                       (reduce_scatter)
//   2. applies Adadelta / SGD-momentum / Adam(W) to its local copy of the parameters and its state shard (update)
//   3. pushes the new parameter values into every peer's parameter buffer (NVLink stores)        (all_gather)
// so the transfer of one vector overlaps the math of the others and nothing is staged.  One barrier before (all
// gradients written) and one after (all pushes landed); epochs on the CH_OPT signal channel.  With NVLS the pull
// becomes multimem.ld_reduce and the push multimem.st.  The gradient copies of a thread's NEXT vector are requested before
// the current one is processed, so two NVLink round trips per thread overlap the update math and the stores.
#include "coll/coll_launch.cuh"
#include "ops/optim_math.cuh"

namespace {
constexpr int kThreads = kCollThreads;

template <int KIND, typename G, bool kNvls>
struct FusedRsOptAgBody {
  static constexpr int kThreads = kCollThreads;
  struct Args { size_t goff; size_t poff; float* s1; float* s2; size_t n; const float* hp; int chan; };
  // the W copies (or their in-switch sum) of gradient vector v
  static __device__ __forceinline__ void fetch(const DcpDevCtx& c, size_t goff, size_t v, uint4 (&raw)[DCP_MAX_RANKS]) {
    if constexpr (kNvls && (std::is_same<G, float>::value || std::is_same<G, __nv_bfloat16>::value)) {
      raw[0] = multimem_ld_reduce_add<G>(c.mc + goff + v * 16);
    } else {
#pragma unroll
      for (int p = 0; p < DCP_MAX_RANKS; ++p) if (p < c.world) raw[p] = ld16_stream(c.peer[p] + goff + v * 16);
    }
  }
  static __device__ __forceinline__ void run(const DcpDevCtx& c, const Args& a) {
  const size_t goff = a.goff, poff = a.poff, n = a.n;
  float* __restrict__ s1 = a.s1; float* __restrict__ s2 = a.s2;
  const float* __restrict__ hp = a.hp;
  const int chan = a.chan;
  constexpr int GV = 16 / sizeof(G);           // gradient elements per 16-byte vector
  const int W = c.world;
  const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (size_t)gridDim.x * blockDim.x;
  // shard boundaries in units of GV elements so that gradient vectors never straddle two ranks
  const size_t nvec = (n + GV - 1) / GV;
  const size_t per = (nvec + W - 1) / W;
  const size_t vb = min(per * c.rank, nvec), ve = min(vb + per, nvec);
  const OptHyper H = load_hyper<KIND>(hp);
  const float gs = hp[7], inv_w = 1.f / (float)W;
  if (!dcp_barrier(c, chan)) return;
  uint4 raw[DCP_MAX_RANKS], nxt[DCP_MAX_RANKS];
  if (vb + tid < ve) fetch(c, goff, vb + tid, raw);
  for (size_t v = vb + tid; v < ve; v += stride) {
    if (v + stride < ve) fetch(c, goff, v + stride, nxt);      // next vector's round trip starts now
    float g[GV];
    if constexpr (kNvls && (std::is_same<G, float>::value || std::is_same<G, __nv_bfloat16>::value)) {
      Vec<G> x; x.u = raw[0];
#pragma unroll
      for (int i = 0; i < GV; ++i) g[i] = ElemTraits<G>::up(x.e[i]) * inv_w;
    } else {
#pragma unroll
      for (int i = 0; i < GV; ++i) g[i] = 0.f;
#pragma unroll
      for (int p = 0; p < DCP_MAX_RANKS; ++p) if (p < W) {
        Vec<G> x; x.u = raw[p];
#pragma unroll
        for (int i = 0; i < GV; ++i) g[i] += ElemTraits<G>::up(x.e[i]);
      }
#pragma unroll
      for (int i = 0; i < GV; ++i) g[i] *= inv_w;
    }
    // parameters are fp32: GV of them = GV/4 vectors
    const size_t e0 = v * GV;                       // first element index
    float* mine = reinterpret_cast<float*>(c.peer[c.rank] + poff);
    float pv[GV];
#pragma unroll
    for (int i = 0; i < GV; i += 4) *reinterpret_cast<float4*>(&pv[i]) = *reinterpret_cast<const float4*>(mine + e0 + i);
    const size_t sidx = e0 - vb * GV;               // index into the local state shard (a multiple of GV: 16-byte aligned)
    if (e0 + GV <= n) {
      // whole vector inside the tensor (all but possibly the last one): 16-byte state accesses
      float a[GV], b[GV];
#pragma unroll
      for (int i = 0; i < GV; i += 4) {
        *reinterpret_cast<float4*>(&a[i]) = s1 ? *reinterpret_cast<const float4*>(s1 + sidx + i) : make_float4(0.f, 0.f, 0.f, 0.f);
        *reinterpret_cast<float4*>(&b[i]) = s2 ? *reinterpret_cast<const float4*>(s2 + sidx + i) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
#pragma unroll
      for (int i = 0; i < GV; ++i) opt_update_h<KIND>(pv[i], g[i] * gs, a[i], b[i], H);
#pragma unroll
      for (int i = 0; i < GV; i += 4) {
        if (s1) *reinterpret_cast<float4*>(s1 + sidx + i) = *reinterpret_cast<const float4*>(&a[i]);
        if (s2) *reinterpret_cast<float4*>(s2 + sidx + i) = *reinterpret_cast<const float4*>(&b[i]);
      }
    } else {
#pragma unroll
      for (int i = 0; i < GV; ++i) {
        if (e0 + i < n) {
          float a = s1 ? s1[sidx + i] : 0.f, b = s2 ? s2[sidx + i] : 0.f;
          opt_update_h<KIND>(pv[i], g[i] * gs, a, b, H);
          if (s1) s1[sidx + i] = a;
          if (s2) s2[sidx + i] = b;
        }
      }
    }
#pragma unroll
    for (int i = 0; i < GV; i += 4) {
      const uint4 out = *reinterpret_cast<const uint4*>(&pv[i]);
      if constexpr (kNvls) {
        multimem_st16(c.mc + poff + (e0 + i) * 4, out);
      } else {
#pragma unroll
        for (int p = 0; p < DCP_MAX_RANKS; ++p) if (p < W) st16_stream(c.peer[p] + poff + (e0 + i) * 4, out);
      }
    }
#pragma unroll
    for (int p = 0; p < DCP_MAX_RANKS; ++p) raw[p] = nxt[p];
  }
  if (!dcp_barrier(c, chan)) return;
  }
};
}  // namespace

extern "C" {
const DcpDevCtx* dcp_symm_ctx(void* h);
size_t dcp_symm_bytes(void* h);

// Elements of the shard rank `rank` owns for a bucket of n elements of `grad_dtype` (state tensors need this many floats).
size_t dcp_fused_shard_elems(size_t n, int world, int grad_dtype) {
  const size_t gv = grad_dtype == DCP_BF16 ? 8 : 4;
  const size_t nvec = (n + gv - 1) / gv;
  return (nvec + world - 1) / world * gv;
}

// grads / params: device pointers INSIDE the symmetric heap (same offsets on every rank); the parameter buffer must
// be padded to a multiple of 8 floats.  s1 / s2: this rank's state shards (dcp_fused_shard_elems floats, or null).
// kind: 0 Adadelta, 1 SGD, 2 Adam.  hp: device hyper-parameter block (see ops/optim.cu; call dcp_optim_advance first).
int dcp_fused_rs_opt_ag(void* h, const void* grads, float* params, float* s1, float* s2, size_t n, const float* hp, int kind, int grad_dtype,
                        int use_nvls, int chan, int max_ctas, cudaStream_t stream) {
  const DcpDevCtx& c = *dcp_symm_ctx(h);
  const char* base = c.peer[c.rank];
  const size_t heap = dcp_symm_bytes(h);
  const char* g = static_cast<const char*>(grads); const char* p = reinterpret_cast<const char*>(params);
  if (g < base || g >= base + heap || p < base || p >= base + heap) { dcp_set_error("fused step: gradients and parameters must live in the symmetric heap"); return -1; }
  const size_t goff = g - base, poff = p - base;
  if ((goff | poff) & 15) { dcp_set_error("fused step: buffers must be 16-byte aligned"); return -1; }
  if (grad_dtype != DCP_F32 && grad_dtype != DCP_BF16) { dcp_set_error("fused step: gradients must be f32 or bf16"); return -1; }
  const bool nvls = use_nvls && c.nvls;
  const size_t shard_bytes = n * 4 / (c.world ? c.world : 1);
  long grid = (long)((shard_bytes + (size_t)kThreads * 16 * 4 - 1) / ((size_t)kThreads * 16 * 4));
  if (grid < 1) grid = 1;
  if (max_ctas <= 0) max_ctas = 148;        // the shard is streamed once (5-7 arrays): every SM has to pull (measured: 32 CTAs 753 us, 148 222 us, 296 238 us for 32 M elements)
  if (grid > max_ctas) grid = max_ctas;
  if (grid > DCP_MAX_BLOCKS) grid = DCP_MAX_BLOCKS;
#define LAUNCH(K, G, NV) coll_launch<FusedRsOptAgBody<K, G, NV>>(h, c, {goff, poff, s1, s2, n, hp, chan}, (int)grid, stream)
#define BY_KIND(G, NV) do { if (kind == 0) LAUNCH(0, G, NV); else if (kind == 1) LAUNCH(1, G, NV); else LAUNCH(2, G, NV); } while (0)
  if (grad_dtype == DCP_F32) { if (nvls) BY_KIND(float, true); else BY_KIND(float, false); }
  else { if (nvls) BY_KIND(__nv_bfloat16, true); else BY_KIND(__nv_bfloat16, false); }
#undef BY_KIND
#undef LAUNCH
  DCP_CUDA_CHECK(cudaGetLastError());
  return 0;
}
}  // extern "C"
