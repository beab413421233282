// Data-parallel tail of the flagship training step, fused into the kernels around it (ops/dp_push.cuh has the protocol).
// In the reference this is torch's C++ reducer launching an all-reduce per gradient bucket behind loss.backward()
// and optimizer.step() afterwards (ref:main.py:62-63,122); here the last gradient bucket never sees a collective launch:
//
//   k_splitk_reduce_push   finishes conv2's split-K weight / bias gradient and pushes the sums straight into every rank's
//                          payload block (multimem.st through NVSwitch, or W peer stores), last CTA signals
//   k_conv1_wgrad (dp.on)  same for conv1's 320 gradients (convnet_fused.cu)
//   k_dp_tail              waits for all producers of all ranks, adds the W slots, averages, writes the gradients and applies
//                          the optimizer update (Adadelta / SGD / Adam, ops/optim_math.cuh) in the same pass
#include "ops/dp_push.cuh"
#include "ops/optim_math.cuh"

extern "C" const DcpDevCtx* dcp_symm_ctx(void* h);
extern "C" size_t dcp_symm_bytes(void* h);

int dcp_dp_nvls_mode(int set) {
  static int mode = [] { const char* e = getenv("DCP_DP_NVLS"); return (e && e[0] == '1') ? 1 : 0; }();
  if (set >= 0) mode = set ? 1 : 0;
  return mode;
}
// Payload path of the fused data-parallel tail: 1 = multimem.st through NVSwitch, 0 = one store per peer (default).
extern "C" int dcp_dp_set_nvls(int on) { dcp_dp_nvls_mode(on ? 1 : 0); return 0; }

namespace {

// out[v] = sum_z scratch[z][v] for the n / 4 float4 columns; CTA = 32 columns x 8 split lanes (as k_splitk_reduce in gemm_tc.cu).
// Column v goes to float index 4 v of this producer's range in the payload slot.
__global__ void __launch_bounds__(256) k_splitk_reduce_push(const float* __restrict__ scratch, int nsplit, long n, const DpPush dp) {
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
    dp_push4(dp, dp_slot_off(dp, (unsigned)(*dp.step) & 1u), (size_t)v * 4, acc);
  }
  dp_finish_cta(dp, gridDim.x);
}

struct DpTail {
  int nsig, chan;
  int blks[4];              // signal words of the producers: dcp_sig(own heap, chan, blks[i], source rank)
  int row_sig[16];          // producer (index into blks) that table row i depends on
  int pull;                 // slot p is read from rank p's heap (producers wrote only their own copy)
  size_t pay_off;
  unsigned slot_elems;
  const long long* step;
  const float* bucket;      // first float of the gradient bucket: entry.g - bucket = float index inside a slot
  DcpDevCtx c;
};

template <int KIND>
__global__ void __launch_bounds__(256) k_dp_tail(const DcpOptEntry* __restrict__ table, const float* __restrict__ hp, const DpTail t) {
  __shared__ int s_ok;
  griddep_launch();
  griddep_wait();
  const int W = t.c.world;
  const uint32_t epoch = (uint32_t)(*t.step);
  char* me = t.c.peer[t.c.rank];
  if (threadIdx.x == 0) s_ok = 1;
  __syncthreads();
  // a CTA only waits for the producer of ITS tensor: rows whose gradients left early (conv2) are updated while the signals of
  // the last producer (conv1) are still in flight
  if ((int)threadIdx.x < W) {
    if (!wait_ge(t.c, dcp_sig(me, t.chan, t.blks[t.row_sig[blockIdx.y]], threadIdx.x), epoch)) s_ok = 0;
  }
  __syncthreads();
  if (!s_ok) return;        // a peer never arrived: the host sees the error flag; do not touch parameters with partial sums
  const DcpOptEntry e = table[blockIdx.y];
  const OptHyper H = load_hyper<KIND>(hp);
  const float gs = hp[7] / (float)W;
  float* g = const_cast<float*>(static_cast<const float*>(e.g));
  const size_t goff = (size_t)(g - t.bucket);
  // slot p: in this rank's block (pushed by rank p) or, in pull mode, in rank p's own block
  const float* slot[DCP_MAX_RANKS];
#pragma unroll
  for (int p = 0; p < DCP_MAX_RANKS; ++p)
    if (p < W) slot[p] = reinterpret_cast<const float*>((t.pull ? t.c.peer[p] : me) + t.pay_off) + ((size_t)(epoch & 1u) * W + p) * t.slot_elems + goff;
  const float* slots = slot[0];
  const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x, nthr = (long long)gridDim.x * blockDim.x;
  const bool vec = (((uintptr_t)e.p | (uintptr_t)e.s1 | (uintptr_t)e.s2 | (uintptr_t)g | (uintptr_t)slots) & 15) == 0;
  const long long n4 = vec ? e.n / 4 : 0;
  for (long long i = tid; i < n4; i += nthr) {
    uint4 raw[DCP_MAX_RANKS];
#pragma unroll
    for (int p = 0; p < DCP_MAX_RANKS; ++p) if (p < W) raw[p] = ld16_stream(reinterpret_cast<const float4*>(slot[p]) + i);   // all W loads in flight
    float4 s4 = make_float4(__uint_as_float(raw[0].x), __uint_as_float(raw[0].y), __uint_as_float(raw[0].z), __uint_as_float(raw[0].w));
#pragma unroll
    for (int p = 1; p < DCP_MAX_RANKS; ++p) if (p < W) {
      s4.x += __uint_as_float(raw[p].x); s4.y += __uint_as_float(raw[p].y); s4.z += __uint_as_float(raw[p].z); s4.w += __uint_as_float(raw[p].w);
    }
    float4 p4 = reinterpret_cast<float4*>(e.p)[i];
    float4 a4 = e.s1 ? reinterpret_cast<float4*>(e.s1)[i] : make_float4(0.f, 0.f, 0.f, 0.f);
    float4 b4 = e.s2 ? reinterpret_cast<float4*>(e.s2)[i] : make_float4(0.f, 0.f, 0.f, 0.f);
    float* gv = reinterpret_cast<float*>(&s4);
    float* pp = reinterpret_cast<float*>(&p4); float* pa = reinterpret_cast<float*>(&a4); float* pb = reinterpret_cast<float*>(&b4);
#pragma unroll
    for (int j = 0; j < 4; ++j) { gv[j] *= gs; opt_update_h<KIND>(pp[j], gv[j], pa[j], pb[j], H); }
    reinterpret_cast<float4*>(g)[i] = s4;            // .grad holds the averaged gradient, as after DDP's all-reduce
    reinterpret_cast<float4*>(e.p)[i] = p4;
    if (e.s1) reinterpret_cast<float4*>(e.s1)[i] = a4;
    if (e.s2) reinterpret_cast<float4*>(e.s2)[i] = b4;
  }
  for (long long i = n4 * 4 + tid; i < e.n; i += nthr) {
    float gsum = __ldcg(slots + i);
    for (int p = 1; p < W; ++p) gsum += __ldcg(slot[p] + i);
    gsum *= gs;
    float p = e.p[i], s1 = e.s1 ? e.s1[i] : 0.f, s2 = e.s2 ? e.s2[i] : 0.f;
    opt_update_h<KIND>(p, gsum, s1, s2, H);
    g[i] = gsum; e.p[i] = p;
    if (e.s1) e.s1[i] = s1;
    if (e.s2) e.s2[i] = s2;
  }
}

int check_payload(const DcpDevCtx& c, void* h, const void* payload, unsigned slot_elems, const char* who) {
  const char* base = c.peer[c.rank];
  const char* pay = static_cast<const char*>(payload);
  if (pay < base || pay + (size_t)2 * c.world * slot_elems * 4 > base + dcp_symm_bytes(h) || ((pay - base) & 15) || (slot_elems & 3)) {
    dcp_set_error("%s: the payload block must be a 16-byte aligned region of the symmetric heap holding 2 * world * slot_elems floats", who);
    return -1;
  }
  return 0;
}

}  // namespace

extern "C" {

// Finish a split-K reduction (scratch: [nsplit][n] floats, n % 4 == 0) and push the n sums into floats
// [elem_off, elem_off + n) of this rank's slot in every rank's payload block; the last CTA signals (chan, blk).
// Nothing is written locally: dcp_dp_tail produces the averaged gradients.  ticket: one zeroed u32 in device memory
// owned by the caller.  step: device int64 counter that holds the same value on every rank when the kernel runs.
int dcp_splitk_reduce_push(void* h, const float* scratch, int nsplit, long n, void* payload, unsigned slot_elems, unsigned elem_off,
                           int chan, int blk, const void* step, void* ticket, int pull, cudaStream_t s) {
  const DcpDevCtx& c = *dcp_symm_ctx(h);
  if (check_payload(c, h, payload, slot_elems, "splitk reduce push")) return -1;
  if ((n & 3) || (elem_off & 3) || elem_off + n > slot_elems || chan < 0 || chan >= DCP_NUM_CHANNELS || blk < 0 || blk >= DCP_MAX_BLOCKS || !step || !ticket) {
    dcp_set_error("splitk reduce push: bad arguments"); return -1;
  }
  DpPush dp{};
  dp.on = 1; dp.chan = chan; dp.blk = blk; dp.nvls = dp_use_nvls(c);
  dp.pay_off = (size_t)(static_cast<const char*>(payload) - c.peer[c.rank]); dp.slot_elems = slot_elems; dp.elem_off = elem_off;
  dp.step = static_cast<const long long*>(step); dp.ticket = static_cast<unsigned*>(ticket); dp.c = c; dp.pull = pull ? 1 : 0;
  DCP_CUDA_CHECK(dcp_launch_pdl(k_splitk_reduce_push, dim3((unsigned)((n / 4 + 31) / 32)), dim3(256), 0, s, scratch, nsplit, n, dp));
  return 0;
}

// Consumer of the pushed gradients: for the `ntensors` rows of the optimizer table (param, grad, state1, state2, numel; every
// grad a view into the bucket that starts at `bucket`, in bucket order = slot order) wait for the `nsig` producers
// (signal words blks[i] on `chan`) of every rank, average the W slots, store the averaged gradient and apply the update.
// row_sig[i] (i < ntensors <= 16) names the producer (index into blks) whose signal guards row i's gradients.
// kind: 0 Adadelta, 1 SGD, 2 Adam; hp: the optimizer's device hyper-parameter block (dcp_optim_advance'd by the caller).
int dcp_dp_tail(void* h, const void* table, int ntensors, long long max_numel, const float* hp, int kind, const void* bucket,
                void* payload, unsigned slot_elems, int chan, const int* blks, int nsig, const int* row_sig, const void* step, int pull, cudaStream_t s) {
  const DcpDevCtx& c = *dcp_symm_ctx(h);
  if (check_payload(c, h, payload, slot_elems, "dp tail")) return -1;
  if (ntensors <= 0 || ntensors > 16 || nsig < 1 || nsig > 4 || chan < 0 || chan >= DCP_NUM_CHANNELS || !step || !bucket || !row_sig) {
    dcp_set_error("dp tail: bad arguments"); return -1;
  }
  DpTail t{};
  t.nsig = nsig; t.chan = chan; t.pull = pull ? 1 : 0;
  for (int i = 0; i < nsig; ++i) {
    if (blks[i] < 0 || blks[i] >= DCP_MAX_BLOCKS) { dcp_set_error("dp tail: bad signal slot %d", blks[i]); return -1; }
    t.blks[i] = blks[i];
  }
  for (int i = 0; i < ntensors; ++i) {
    if (row_sig[i] < 0 || row_sig[i] >= nsig) { dcp_set_error("dp tail: row %d names producer %d of %d", i, row_sig[i], nsig); return -1; }
    t.row_sig[i] = row_sig[i];
  }
  t.pay_off = (size_t)(static_cast<const char*>(payload) - c.peer[c.rank]); t.slot_elems = slot_elems;
  t.step = static_cast<const long long*>(step); t.bucket = static_cast<const float*>(bucket); t.c = c;
  int gx = (int)((max_numel + 256 * 4 - 1) / (256 * 4)); if (gx < 1) gx = 1; if (gx > 148) gx = 148;
  dim3 grid(gx, ntensors);
  const DcpOptEntry* tb = static_cast<const DcpOptEntry*>(table);
  if (kind == 0) DCP_CUDA_CHECK(dcp_launch_pdl(k_dp_tail<0>, grid, dim3(256), 0, s, tb, hp, t));
  else if (kind == 1) DCP_CUDA_CHECK(dcp_launch_pdl(k_dp_tail<1>, grid, dim3(256), 0, s, tb, hp, t));
  else DCP_CUDA_CHECK(dcp_launch_pdl(k_dp_tail<2>, grid, dim3(256), 0, s, tb, hp, t));
  return 0;
}

}  // extern "C"
