// Shared definitions for libdcp_cuda.so (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdio>

#define DCP_MAX_RANKS 8
#define DCP_NUM_CHANNELS 8
#define DCP_MAX_BLOCKS 512       // upper bound on the grid of any collective kernel
#define DCP_P2P_BLOCKS 128       // largest grid of the send/recv kernels (one flag / ack word per block and peer)

// dtype codes shared with Python (ops/cuda.py, parallel/p2p_backend.py)
enum DcpDtype : int { DCP_F32 = 0, DCP_BF16 = 1, DCP_F16 = 2, DCP_I32 = 3, DCP_I64 = 4, DCP_F64 = 5, DCP_U8 = 6 };
// reduce ops (same numbering as types.ReduceOp)
enum DcpOp : int { DCP_SUM = 0, DCP_PROD = 1, DCP_MIN = 2, DCP_MAX = 3, DCP_BAND = 4, DCP_BOR = 5, DCP_BXOR = 6, DCP_AVG = 7 };

// Thread-local last error string for the C ABI.
extern "C" const char* dcp_last_error();
void dcp_set_error(const char* fmt, ...);

#define DCP_CUDA_CHECK(expr)                                                                       \
  do {                                                                                             \
    cudaError_t _e = (expr);                                                                       \
    if (_e != cudaSuccess) {                                                                       \
      dcp_set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__);   \
      return -1;                                                                                   \
    }                                                                                              \
  } while (0)

// ------------------------------------------------------------------------------------------------
// Programmatic dependent launch (PDL).  A kernel launched with dcp_launch_pdl may become resident while its
// predecessors in the stream are still draining: its prologue (barrier init, TMEM allocation, descriptor prefetch)
// then overlaps their tails and the launch latency disappears.  The contract inside such a kernel:
//   * griddep_launch() early lets the NEXT kernel in the stream start its own prologue;
//   * griddep_wait() must come before the first access to global memory that any earlier kernel wrote or will
//     still read - it returns once every prerequisite grid has completed and its writes are visible.
// Launched without the attribute (DCP_PDL=0, or after a kernel that knows nothing of this) both are no-ops and the
// stream order is the usual full serialisation.
// ------------------------------------------------------------------------------------------------
#ifdef __CUDACC__
__device__ __forceinline__ void griddep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void griddep_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
bool dcp_pdl_enabled();
template <typename... P, typename... A>
inline cudaError_t dcp_launch_pdl(void (*kern)(P...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, A... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at; cfg.numAttrs = dcp_pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<P>(args)...);
}
#endif

// ------------------------------------------------------------------------------------------------
// Symmetric heap layout (identical offsets on every rank).  All sizes in bytes.
// ------------------------------------------------------------------------------------------------
struct DcpHeapLayout {
  // written by peers
  static constexpr size_t kSigOff = 0;                                                   // u32 [chan][block][src]
  static constexpr size_t kSigBytes = (size_t)DCP_NUM_CHANNELS * DCP_MAX_BLOCKS * DCP_MAX_RANKS * 4;
  static constexpr size_t kP2PFlagOff = kSigOff + kSigBytes;                             // u32 [src][block]
  static constexpr size_t kP2PFlagBytes = (size_t)DCP_MAX_RANKS * DCP_P2P_BLOCKS * 4;
  static constexpr size_t kP2PAckOff = kP2PFlagOff + kP2PFlagBytes;                      // u32 [dst][block]
  static constexpr size_t kP2PAckBytes = kP2PFlagBytes;
  // local only
  static constexpr size_t kEpochOff = kP2PAckOff + kP2PAckBytes;                         // u32 [chan][block]
  static constexpr size_t kEpochBytes = (size_t)DCP_NUM_CHANNELS * DCP_MAX_BLOCKS * 4;
  static constexpr size_t kCallsOff = kEpochOff + kEpochBytes;                           // u32 [chan][block]
  static constexpr size_t kCallsBytes = kEpochBytes;
  static constexpr size_t kSendSeqOff = kCallsOff + kCallsBytes;                         // u32 [dst][block]
  static constexpr size_t kRecvSeqOff = kSendSeqOff + kP2PFlagBytes;                     // u32 [src][block]
  static constexpr size_t kGCallsOff = kRecvSeqOff + kP2PFlagBytes;                      // u32 [chan][2]: launches completed, CTA ticket
  static constexpr size_t kGCallsBytes = (size_t)DCP_NUM_CHANNELS * 2 * 4;
  static constexpr size_t kCtlBytes = 1u << 20;                                          // control region, zeroed at init
  static_assert(kGCallsOff + kGCallsBytes <= kCtlBytes, "control region overflow");
};

// Passed by value to every collective kernel.
struct DcpDevCtx {
  int rank, world;
  int nvls;             // multicast mapping available
  int p2p_blocks;       // grid limit of send / recv on this heap (<= DCP_P2P_BLOCKS; an emulated world keeps it small)
  char* peer[DCP_MAX_RANKS];   // peer[r] = VA (in this process) of rank r's heap; peer[rank] = own heap
  char* mc;                    // multicast VA of the heap (NVLS) or nullptr
  unsigned long long timeout_ns;
  int* err;                    // host-mapped flag: set non-zero when a wait timed out
  size_t p2p_slot_off, p2p_slot_bytes;   // receive slots: [src] x slot_bytes
  size_t staging_off, staging_bytes;     // bounce buffer for non-symmetric tensors
};

// ------------------------------------------------------------------------------------------------
// Device helpers: system-scope release/acquire on signal words, bounded spins.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// Spin until *p >= target (wrap-safe).  Returns false after ctx.timeout_ns and raises ctx.err: a dead
// peer becomes an error on the host instead of a hung GPU.
__device__ __forceinline__ bool wait_ge(const DcpDevCtx& c, const uint32_t* p, uint32_t target) {
  if ((int32_t)(ld_acquire_sys(p) - target) >= 0) return true;
  unsigned long long t0 = globaltimer_ns();
  for (;;) {
#pragma unroll 1
    for (int i = 0; i < 64; ++i)
      if ((int32_t)(ld_acquire_sys(p) - target) >= 0) return true;
    if (globaltimer_ns() - t0 > c.timeout_ns) {
      if (c.err) *((volatile int*)c.err) = 1;
      return false;
    }
  }
}

__device__ __forceinline__ uint32_t* dcp_sig(char* heap, int chan, int block, int src) {
  return reinterpret_cast<uint32_t*>(heap + DcpHeapLayout::kSigOff) + ((size_t)chan * DCP_MAX_BLOCKS + block) * DCP_MAX_RANKS + src;
}
__device__ __forceinline__ uint32_t* dcp_epoch(char* heap, int chan, int block) {
  return reinterpret_cast<uint32_t*>(heap + DcpHeapLayout::kEpochOff) + (size_t)chan * DCP_MAX_BLOCKS + block;
}
__device__ __forceinline__ uint32_t* dcp_calls(char* heap, int chan, int block) {
  return reinterpret_cast<uint32_t*>(heap + DcpHeapLayout::kCallsOff) + (size_t)chan * DCP_MAX_BLOCKS + block;
}

// Launch counter of a channel (local control memory): the same value in every CTA of a launch and - because all ranks
// issue the same sequence of collectives - on every rank.  Staged kernels whose copy-in and reduce partitions differ pick
// their bounce half from its parity.  dcp_gcalls_finish() is called by every CTA as its last action: the CTA that draws the
// last ticket advances the counter, i.e. only after every CTA of the launch has read it.
__device__ __forceinline__ uint32_t dcp_gcalls(char* heap, int chan) {
  return *((volatile uint32_t*)(heap + DcpHeapLayout::kGCallsOff) + chan * 2);
}
__device__ __forceinline__ void dcp_gcalls_finish(char* heap, int chan) {
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t* w = reinterpret_cast<uint32_t*>(heap + DcpHeapLayout::kGCallsOff) + chan * 2;
    if (atomicAdd(w + 1, 1u) == gridDim.x - 1) { w[1] = 0; __threadfence(); *((volatile uint32_t*)w) = *((volatile uint32_t*)w) + 1; }
  }
}

// Barrier between the CTAs with this blockIdx.x on every rank.  All threads of the CTA must call it.
// Release/acquire at system scope: writes issued by any thread of this CTA before the call are
// visible to the peer CTAs after they return, and vice versa.  Epochs increase monotonically, so the
// signal words never need resetting and the barrier is CUDA-graph replayable.
// Returns false (to every thread of the CTA) when a peer did not arrive within ctx.timeout_ns: the caller then skips its data
// phase instead of reducing / copying whatever happens to be in the buffers (the error flag is already raised for the host).
__device__ __forceinline__ bool dcp_barrier(const DcpDevCtx& c, int chan) {
  __syncthreads();
  char* me = c.peer[c.rank];
  uint32_t* ep = dcp_epoch(me, chan, blockIdx.x);
  const uint32_t e = *((volatile uint32_t*)ep) + 1;
  int ok = 1;
  if (threadIdx.x < c.world) {
    const int peer = threadIdx.x;
    // the release store is cumulative over everything the CTA wrote before the __syncthreads() above: no separate fence
    // (st.release.sys already costs one system-scope drain; a second __threadfence_system() in front doubled it)
    st_release_sys(dcp_sig(c.peer[peer], chan, blockIdx.x, c.rank), e);
    ok = wait_ge(c, dcp_sig(me, chan, blockIdx.x, peer), e) ? 1 : 0;
  }
  ok = __syncthreads_and(ok);
  if (threadIdx.x == 0) *((volatile uint32_t*)ep) = e;
  return ok != 0;
}

// Signal-only / wait-only halves (used when the payload was pushed before the signal).
__device__ __forceinline__ uint32_t dcp_signal_all(const DcpDevCtx& c, int chan) {
  __syncthreads();
  char* me = c.peer[c.rank];
  const uint32_t e = *((volatile uint32_t*)dcp_epoch(me, chan, blockIdx.x)) + 1;
  if (threadIdx.x < c.world) st_release_sys(dcp_sig(c.peer[threadIdx.x], chan, blockIdx.x, c.rank), e);
  return e;
}
__device__ __forceinline__ bool dcp_wait_all(const DcpDevCtx& c, int chan, uint32_t e) {
  char* me = c.peer[c.rank];
  int ok = 1;
  if (threadIdx.x < c.world) ok = wait_ge(c, dcp_sig(me, chan, blockIdx.x, threadIdx.x), e) ? 1 : 0;
  ok = __syncthreads_and(ok);
  if (threadIdx.x == 0) *((volatile uint32_t*)dcp_epoch(me, chan, blockIdx.x)) = e;
  return ok != 0;       // false: a peer never signalled - skip the data phase (see dcp_barrier)
}
