// One-shot gradient exchange fused into the kernels that PRODUCE the gradients (data-parallel tail of the
// flagship step; what torch's reducer does with a separate all-reduce per bucket behind loss.backward(),
// ref:main.py:62,122).
//
// Every rank owns a payload block at the same offset of its symmetric heap: [2 halves][W ranks][slot_elems] floats.
// A producer kernel on rank r writes its results into slot [step & 1][r] of EVERY rank's block (one multimem.st per
// 16 bytes when the NVLS mapping exists, W peer stores otherwise) and, when its last CTA is done, stores the step
// number into one signal word per peer.  Nobody waits in the producers.  The consumer (k_dp_tail, convnet_dp.cu) runs
// at the very end of the step: it waits for the signal words of all producers of all ranks, adds the W slots up,
// averages, and applies the optimizer update in the same pass - reduce + update in one kernel, no separate
// all-reduce launch, no stream hop, and the NVLink round trip overlaps whatever the stream runs in between.
//
// Why two halves are enough: rank A can only push step s + 2 after its own consumer of step s + 1 has seen every
// peer's step s + 1 signals, and a peer sends those after its consumer of step s (the last reader of that half) has
// finished - stream order on the peer.  The step number comes from a device counter that all ranks advance in lock
// step, so the protocol needs no per-call bookkeeping and replays inside a CUDA graph.
#pragma once
#include <cstdlib>

#include "coll/coll_common.cuh"

struct DpPush {
  int on;                   // 0: single rank, plain behaviour
  int chan, blk;            // signal word of this producer: dcp_sig(heap, chan, blk, source rank)
  int nvls;                 // payload through multimem.st
  int pull;                 // 1: write ONLY this rank's own copy of its slot; the consumer reads the peers' copies over NVLink.
                            // No peer stores, hence no system-scope drain per CTA (measured: ~10 us for a 145-CTA producer);
                            // the price is one round of remote loads in the consumer.
  size_t pay_off;           // heap offset of the payload block (identical on every rank, 16-byte aligned)
  unsigned slot_elems;      // floats per (half, rank) slot, multiple of 4
  unsigned elem_off;        // first float of this producer inside a slot, multiple of 4
  const long long* step;    // device step counter (same value on every rank when the kernel runs)
  unsigned* ticket;         // last-CTA ticket of this producer (zero between launches); unused by single-CTA producers
  DcpDevCtx c;
};

// Payload path.  The pushed blocks are small (the flagship's last bucket is 75 KB): plain peer stores have the lower latency;
// multimem.st (one store replicated by the switch) only pays when the per-rank egress matters.  DCP_DP_NVLS=1 selects it.
int dcp_dp_nvls_mode(int set);   // convnet_dp.cu: -1 = query, 0 / 1 = set (dcp_dp_set_nvls); initial value from DCP_DP_NVLS
inline int dp_use_nvls(const DcpDevCtx& c) { return dcp_dp_nvls_mode(-1) && c.nvls && c.mc != nullptr; }

__device__ __forceinline__ size_t dp_slot_off(const DpPush& d, unsigned half) {
  return d.pay_off + (((size_t)half * d.c.world + d.c.rank) * d.slot_elems + d.elem_off) * sizeof(float);
}
// 16 bytes at float index `idx` (multiple of 4) of this producer's range, to every rank (own copy included)
__device__ __forceinline__ void dp_push4(const DpPush& d, size_t slot_off, size_t idx, float4 v) {
  const uint4 u = make_uint4(__float_as_uint(v.x), __float_as_uint(v.y), __float_as_uint(v.z), __float_as_uint(v.w));
  const size_t off = slot_off + idx * sizeof(float);
  if (d.pull) {
    st16(d.c.peer[d.c.rank] + off, u);
  } else if (d.nvls) {
    multimem_st16(d.c.mc + off, u);
  } else {
    for (int p = 0; p < d.c.world; ++p) st16_stream(d.c.peer[p] + off, u);
  }
}
__device__ __forceinline__ void dp_push1(const DpPush& d, size_t slot_off, size_t idx, float v) {
  for (int p = 0; p < d.c.world; ++p) reinterpret_cast<float*>(d.c.peer[p] + slot_off)[idx] = v;
}
// Signal every rank that this producer's payload of the current step is complete.  One CTA calls it, after a
// __syncthreads() that follows its last payload store (and after it has established that it is the last CTA).
__device__ __forceinline__ void dp_signal(const DpPush& d) {
  // st.release.sys is cumulative over the CTA's earlier stores (ordered before it by the __syncthreads()): no extra fence
  if ((int)threadIdx.x < d.c.world) st_release_sys(dcp_sig(d.c.peer[threadIdx.x], d.chan, d.blk, d.c.rank), (uint32_t)(*d.step));
}
// Multi-CTA producers: every CTA calls this after its pushes; the CTA that draws the last ticket signals.
__device__ __forceinline__ void dp_finish_cta(const DpPush& d, unsigned num_ctas) {
  __shared__ bool s_dp_last;
  __syncthreads();
  if (threadIdx.x == 0) {
    if (d.pull) __threadfence(); else __threadfence_system();     // local payload: device scope is enough in front of the ticket
    s_dp_last = atomicAdd(d.ticket, 1u) == num_ctas - 1;
    if (s_dp_last) { __threadfence(); *d.ticket = 0; }   // acquire side of the ticket chain: the other CTAs' pushes are ordered before the signal
  }
  __syncthreads();
  if (s_dp_last) dp_signal(d);
}
