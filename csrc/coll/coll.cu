// Hand-written NVLink collectives for sm_100a: every kernel reads / writes peer GPUs' memory directly
// through the symmetric heap (csrc/symm) - no NCCL, no host staging (the reference's gloo backend
// bounces CUDA tensors through pinned host memory, ref:main.py:50; its NCCL baseline uses ncclComm).
//
//   all_reduce  one-shot  : push own copy into every peer's slot, one signal, reduce locally (latency floor)
//               two-shot  : pull-reduce own 1/W slice from all peers, push the result to all peers
//               NVLS      : multimem.ld_reduce (in-switch reduction) + multimem.st (in-switch broadcast)
//   broadcast, all_gather, reduce_scatter, reduce, gather, scatter, send/recv, barrier
//
// Cross-rank ordering uses per-CTA signal words with monotonically increasing epochs (common.cuh),
// so nothing is ever reset and every kernel can be captured in a CUDA graph and replayed.
// Every kernel is a "body" (coll_launch.cuh): the same code runs one rank per GPU in production and all ranks of a small world
// inside one cooperative launch on a single GPU in the emulated test tier.
// Staged (non-symmetric tensor) variants double-buffer the bounce area on the parity of the channel's launch counter
// (common.cuh), which removes the trailing barrier: a rank reuses a half two launches later, after it has passed a barrier
// of the launch in between - so every peer has, in stream order, completed the launch that used that half.
#include <cstdlib>
#include <type_traits>

#include "coll/coll_launch.cuh"

namespace {


struct Span { size_t begin, end; };  // vector index range

__device__ __forceinline__ Span rank_slice(size_t nvec, int world, int r) {
  size_t per = (nvec + world - 1) / world;
  size_t b = per * r; if (b > nvec) b = nvec;
  size_t e = b + per; if (e > nvec) e = nvec;
  return {b, e};
}

// -------------------------------------------------------------------------------------------------
// all_reduce, one-shot push.  slot layout in the one-shot region: [half][src rank][slot_bytes].
// -------------------------------------------------------------------------------------------------
template <typename T>
struct AllreduceOneshotBody {
  static constexpr int kThreads = kCollThreads;
  struct Args { const T* in; T* out; size_t n; float scale; int op; int chan; size_t region_off; size_t slot_bytes; };
  static __device__ __forceinline__ void run(const DcpDevCtx& c, const Args& a) {
    const T* in = a.in;
    T* out = a.out;
    size_t n = a.n;
    float scale = a.scale;
    int op = a.op;
    int chan = a.chan;
    size_t region_off = a.region_off;
    size_t slot_bytes = a.slot_bytes;
  using Tr = ElemTraits<T>; using Acc = typename Tr::Acc;
    constexpr int V = Vec<T>::N;
    const int W = c.world;
    const uint32_t half = (dcp_gcalls(c.peer[c.rank], chan) & 1u);
    const size_t my_slot = region_off + ((size_t)half * W + c.rank) * slot_bytes;
    const bool aligned = ((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(out)) & 15) == 0;
    const size_t nvec = aligned ? n / V : 0;
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (size_t)gridDim.x * blockDim.x;
    for (size_t v = tid; v < nvec; v += stride) {
      uint4 val = ld16(in + v * V);
      for (int p = 0; p < W; ++p) st16_stream(c.peer[p] + my_slot + v * 16, val);
    }
    for (size_t i = nvec * V + tid; i < n; i += stride) {
      T val = in[i];
      for (int p = 0; p < W; ++p) reinterpret_cast<T*>(c.peer[p] + my_slot)[i] = val;
    }
    uint32_t e = dcp_signal_all(c, chan);
    if (!dcp_wait_all(c, chan, e)) return;
    const char* base = c.peer[c.rank] + region_off + (size_t)half * W * slot_bytes;
    for (size_t v = tid; v < nvec; v += stride) {
      Acc acc[V];
      Vec<T> x; x.u = ld16(base + v * 16);
#pragma unroll
      for (int i = 0; i < V; ++i) acc[i] = Tr::up(x.e[i]);
      for (int p = 1; p < W; ++p) {
        x.u = ld16(base + (size_t)p * slot_bytes + v * 16);
#pragma unroll
        for (int i = 0; i < V; ++i) acc[i] = red_apply<Acc>(op, acc[i], Tr::up(x.e[i]));
      }
#pragma unroll
      for (int i = 0; i < V; ++i) x.e[i] = Tr::down(red_finish<T>(acc[i], scale, W, op));
      st16(out + v * V, x.u);
    }
    for (size_t i = nvec * V + tid; i < n; i += stride) {
      Acc acc = Tr::up(reinterpret_cast<const T*>(base)[i]);
      for (int p = 1; p < W; ++p) acc = red_apply<Acc>(op, acc, Tr::up(reinterpret_cast<const T*>(base + (size_t)p * slot_bytes)[i]));
      out[i] = Tr::down(red_finish<T>(acc, scale, W, op));
    }
    dcp_gcalls_finish(c.peer[c.rank], chan);
  }
};

// -------------------------------------------------------------------------------------------------
// all_reduce, two-shot (P2P or NVLS) on a buffer that sits at heap offset `off` on every rank.
// in != nullptr: copy in -> buffer first (staged path); out != nullptr: copy buffer -> out last.
// `n` elements; the buffer must be 16-byte aligned (offsets are); tail elements go to rank 0.
// -------------------------------------------------------------------------------------------------
template <typename T, bool kNvls>
struct AllreduceTwoshotBody {
  static constexpr int kThreads = kCollThreads;
  struct Args { const T* in; T* out; size_t off; size_t half_bytes; size_t n; float scale; int op; int chan; };
  static __device__ __forceinline__ void run(const DcpDevCtx& c, const Args& a) {
    const T* in = a.in;
    T* out = a.out;
    size_t off = a.off;
    size_t half_bytes = a.half_bytes;
    size_t n = a.n;
    float scale = a.scale;
    int op = a.op;
    int chan = a.chan;
  using Tr = ElemTraits<T>; using Acc = typename Tr::Acc;
    constexpr int V = Vec<T>::N;
    const int W = c.world;
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (size_t)gridDim.x * blockDim.x;
    const bool staged = in != nullptr;
    if (staged) off += (size_t)(dcp_gcalls(c.peer[c.rank], chan) & 1u) * half_bytes;
    char* mine = c.peer[c.rank] + off;
    const size_t nvec = n / V;
    if (staged) {
      const bool al = (reinterpret_cast<uintptr_t>(in) & 15) == 0;
      if (al) for (size_t v = tid; v < nvec; v += stride) st16(mine + v * 16, ld16_stream(in + v * V));
      else for (size_t v = tid; v < nvec; v += stride)
        for (int i = 0; i < V; ++i) reinterpret_cast<T*>(mine)[v * V + i] = in[v * V + i];
      for (size_t i = nvec * V + tid; i < n; i += stride) reinterpret_cast<T*>(mine)[i] = in[i];
    }
    if (!dcp_barrier(c, chan)) return;
    const Span s = rank_slice(nvec, W, c.rank);
    // vector v is always handled by the thread with v % stride == tid - in copy-in, reduce and copy-out and
    // on every rank - so the CTA that staged an element (and chose the parity half) is the one that reads it
    const size_t v0 = s.begin + ((tid + stride - (s.begin % stride)) % stride);
    if constexpr (kNvls && NvlsOk<T>::value) {
      char* mc = c.mc + off;
      // four independent in-switch reductions in flight per thread before the (posted) multicast stores: one multimem.ld_reduce
      // is a full round trip through NVSwitch, so the loop is latency-bound unless several overlap - with them the kernel needs
      // a quarter of the CTAs for the same time, which is what lets a gradient bucket hide under the backward pass
      size_t v = v0;
      for (; v + 3 * stride < s.end; v += 4 * stride) {
        uint4 r0 = multimem_ld_reduce_add<T>(mc + v * 16), r1 = multimem_ld_reduce_add<T>(mc + (v + stride) * 16);
        uint4 r2 = multimem_ld_reduce_add<T>(mc + (v + 2 * stride) * 16), r3 = multimem_ld_reduce_add<T>(mc + (v + 3 * stride) * 16);
        if (op == DCP_AVG) { r0 = vec_scale<T>(r0, scale); r1 = vec_scale<T>(r1, scale); r2 = vec_scale<T>(r2, scale); r3 = vec_scale<T>(r3, scale); }
        multimem_st16(mc + v * 16, r0); multimem_st16(mc + (v + stride) * 16, r1);
        multimem_st16(mc + (v + 2 * stride) * 16, r2); multimem_st16(mc + (v + 3 * stride) * 16, r3);
      }
      for (; v < s.end; v += stride) {
        uint4 r = multimem_ld_reduce_add<T>(mc + v * 16);
        if (op == DCP_AVG) r = vec_scale<T>(r, scale);
        multimem_st16(mc + v * 16, r);
      }
    } else {
      for (size_t v = v0; v < s.end; v += stride) {
        Acc acc[V]; Vec<T> x;
        uint4 raw[DCP_MAX_RANKS];
#pragma unroll
        for (int p = 0; p < DCP_MAX_RANKS; ++p) if (p < W) raw[p] = ld16_stream(c.peer[p] + off + v * 16);  // all loads in flight first
        x.u = raw[0];
#pragma unroll
        for (int i = 0; i < V; ++i) acc[i] = Tr::up(x.e[i]);
#pragma unroll
        for (int p = 1; p < DCP_MAX_RANKS; ++p) if (p < W) {
          x.u = raw[p];
#pragma unroll
          for (int i = 0; i < V; ++i) acc[i] = red_apply<Acc>(op, acc[i], Tr::up(x.e[i]));
        }
#pragma unroll
        for (int i = 0; i < V; ++i) x.e[i] = Tr::down(red_finish<T>(acc[i], scale, W, op));
#pragma unroll
        for (int p = 0; p < DCP_MAX_RANKS; ++p) if (p < W) st16_stream(c.peer[p] + off + v * 16, x.u);
      }
    }
    if (c.rank == 0) {  // tail (< one vector) in two-shot fashion as well
      for (size_t i = nvec * V + tid; i < n; i += stride) {
        Acc acc = Tr::up(reinterpret_cast<const T*>(c.peer[0] + off)[i]);
        for (int p = 1; p < W; ++p) acc = red_apply<Acc>(op, acc, Tr::up(reinterpret_cast<const T*>(c.peer[p] + off)[i]));
        T r = Tr::down(red_finish<T>(acc, scale, W, op));
        for (int p = 0; p < W; ++p) reinterpret_cast<T*>(c.peer[p] + off)[i] = r;
      }
    }
    if (!dcp_barrier(c, chan)) return;
    if (out != nullptr) {
      const bool al = (reinterpret_cast<uintptr_t>(out) & 15) == 0;
      if (al) for (size_t v = tid; v < nvec; v += stride) st16(out + v * V, ld16(mine + v * 16));
      else for (size_t v = tid; v < nvec; v += stride)
        for (int i = 0; i < V; ++i) out[v * V + i] = reinterpret_cast<const T*>(mine)[v * V + i];
      for (size_t i = nvec * V + tid; i < n; i += stride) out[i] = reinterpret_cast<const T*>(mine)[i];
    }
    if (staged) dcp_gcalls_finish(c.peer[c.rank], chan);
  }
};

// -------------------------------------------------------------------------------------------------
// Byte movers (dtype-agnostic): broadcast / all_gather / gather / scatter.
// mode bits: which of the W "lanes" this rank pushes and where.
// -------------------------------------------------------------------------------------------------
// The byte range is cut into 16-byte chunks and chunk c is ALWAYS handled by the thread with c % stride == tid,
// whatever the alignment of the two pointers.  The staged kernels rely on this: the CTA that pushed a chunk
// (and chose the parity half from its own call counter) is the CTA that later copies it out.
// bypass_l1: the source is re-written by a peer DURING this kernel (the receive slots of a multi-chunk transfer): read it
// with ld.global.cg so that a line cached by an earlier chunk is never served from L1.
__device__ __forceinline__ uint4 ld16_src(const char* p, bool bypass_l1) { return bypass_l1 ? __ldcg(reinterpret_cast<const uint4*>(p)) : ld16(p); }
__device__ __forceinline__ void copy_bytes(char* dst, const char* src, size_t nbytes, size_t tid, size_t stride, bool stream_store, bool bypass_l1 = false) {
  const uintptr_t mis = reinterpret_cast<uintptr_t>(dst) | reinterpret_cast<uintptr_t>(src);
  const bool al16 = (mis & 15) == 0, al4 = (mis & 3) == 0;
  const size_t nchunk = (nbytes + 15) / 16;
  if (al16) {
    // bulk: 4 independent 16-byte loads in flight per thread (same chunk -> thread mapping as the generic loop)
    const size_t nfull = nbytes / 16;
    size_t c = tid;
    for (; c + 3 * stride < nfull; c += 4 * stride) {
      uint4 x0 = ld16_src(src + c * 16, bypass_l1), x1 = ld16_src(src + (c + stride) * 16, bypass_l1), x2 = ld16_src(src + (c + 2 * stride) * 16, bypass_l1), x3 = ld16_src(src + (c + 3 * stride) * 16, bypass_l1);
      if (stream_store) { st16_stream(dst + c * 16, x0); st16_stream(dst + (c + stride) * 16, x1); st16_stream(dst + (c + 2 * stride) * 16, x2); st16_stream(dst + (c + 3 * stride) * 16, x3); }
      else { st16(dst + c * 16, x0); st16(dst + (c + stride) * 16, x1); st16(dst + (c + 2 * stride) * 16, x2); st16(dst + (c + 3 * stride) * 16, x3); }
    }
    for (; c < nfull; c += stride) { uint4 x = ld16_src(src + c * 16, bypass_l1); if (stream_store) st16_stream(dst + c * 16, x); else st16(dst + c * 16, x); }
    if (nfull < nchunk && (nfull % stride) == tid) for (size_t j = nfull * 16; j < nbytes; ++j) dst[j] = bypass_l1 ? (char)__ldcg(reinterpret_cast<const signed char*>(src + j)) : src[j];
    return;
  }
  for (size_t c = tid; c < nchunk; c += stride) {
    const size_t b = c * 16;
    const size_t len = nbytes - b < 16 ? nbytes - b : 16;
    if (al4 && (len & 3) == 0) {
      for (size_t j = 0; j < len; j += 4) {
        const uint32_t* q = reinterpret_cast<const uint32_t*>(src + b + j);
        *reinterpret_cast<uint32_t*>(dst + b + j) = bypass_l1 ? __ldcg(q) : *q;
      }
    } else {
      for (size_t j = 0; j < len; ++j) dst[b + j] = bypass_l1 ? (char)__ldcg(reinterpret_cast<const signed char*>(src + b + j)) : src[b + j];
    }
  }
}

// One multimem.st per 16 bytes reaches every rank's copy of heap offset `off` (NVSwitch replicates the store);
// the < 16-byte tail goes point to point.  Same chunk -> thread mapping as copy_bytes.
__device__ __forceinline__ void push_all_nvls(const DcpDevCtx& c, size_t off, const char* src, size_t nbytes, size_t tid, size_t stride) {
  const size_t nfull = nbytes / 16;
  const bool al = (reinterpret_cast<uintptr_t>(src) & 15) == 0;
  size_t v = tid;
  if (al) {
    // four independent loads in flight per thread before the (posted) multicast stores: a single source has to fill
    // its whole NVLink egress on its own in broadcast
    for (; v + 3 * stride < nfull; v += 4 * stride) {
      const uint4 x0 = ld16(src + v * 16), x1 = ld16(src + (v + stride) * 16), x2 = ld16(src + (v + 2 * stride) * 16), x3 = ld16(src + (v + 3 * stride) * 16);
      multimem_st16(c.mc + off + v * 16, x0); multimem_st16(c.mc + off + (v + stride) * 16, x1);
      multimem_st16(c.mc + off + (v + 2 * stride) * 16, x2); multimem_st16(c.mc + off + (v + 3 * stride) * 16, x3);
    }
  }
  for (; v < nfull; v += stride) {
    uint4 x;
    if (al) x = ld16(src + v * 16);
    else { unsigned char* b = reinterpret_cast<unsigned char*>(&x); for (int j = 0; j < 16; ++j) b[j] = (unsigned char)src[v * 16 + j]; }
    multimem_st16(c.mc + off + v * 16, x);
  }
  if ((nfull % stride) == tid)
    for (size_t j = nfull * 16; j < nbytes; ++j) { const char x = src[j]; for (int p = 0; p < c.world; ++p) c.peer[p][off + j] = x; }
}

// -------------------------------------------------------------------------------------------------
// all_reduce of an ORDINARY tensor (not in the symmetric heap).  The staging copy is turned into the communication:
//   1. rank r stores piece p of its input into slot [r] of rank p's bounce area (peer stores at link rate; nobody copies
//      its whole input into its own heap first),
//   2. barrier; rank p reduces its W slots from LOCAL memory,
//   3. and stores the reduced slice into the result area of every rank (multimem.st with NVLS, peer stores otherwise),
//   4. barrier; everyone copies the result out of its own heap.
// Every 16-byte vector is pushed, reduced, re-pushed and copied out by the same (CTA, thread) index on the ranks involved, so
// the two per-CTA barriers are all the ordering it needs.  Bounce half from the parity of the channel's launch counter.
// Layout at `off`: [W slots of `per` vectors][result: W * per vectors].
// -------------------------------------------------------------------------------------------------
template <typename T>
struct AllreduceStagedBody {
  static constexpr int kThreads = kCollThreads;
  struct Args { const T* in; T* out; size_t off; size_t half_bytes; size_t n; float scale; int op; int chan; };
  static __device__ __forceinline__ void run(const DcpDevCtx& c, const Args& a) {
    using Tr = ElemTraits<T>; using Acc = typename Tr::Acc;
    constexpr int V = Vec<T>::N;
    const int W = c.world, op = a.op, chan = a.chan;
    const float scale = a.scale;
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (size_t)gridDim.x * blockDim.x;
    char* me = c.peer[c.rank];
    const size_t off = a.off + (size_t)(dcp_gcalls(me, chan) & 1u) * a.half_bytes;
    const size_t nbytes = a.n * sizeof(T), nv = (nbytes + 15) / 16, per = (nv + W - 1) / W;
    const size_t res_off = off + (size_t)W * per * 16;
    const char* in = reinterpret_cast<const char*>(a.in);
    // 1. push
    for (int p = 0; p < W; ++p) {
      const size_t b = per * p < nv ? per * p : nv, e = b + per < nv ? b + per : nv;
      const size_t lo = b * 16, hi = e * 16 < nbytes ? e * 16 : nbytes;
      if (hi > lo) copy_bytes(c.peer[p] + off + (size_t)c.rank * per * 16, in + lo, hi - lo, tid, stride, p != c.rank);
    }
    if (!dcp_barrier(c, chan)) return;
    // 2 + 3. reduce own slice from the local slots, publish it
    {
      const size_t b = per * c.rank < nv ? per * c.rank : nv, e = b + per < nv ? b + per : nv;
      const char* slots = me + off;
      const bool nvls = c.nvls && W > 2;
      for (size_t v = tid; v < e - b; v += stride) {
        Acc acc[V]; Vec<T> x;
        uint4 raw[DCP_MAX_RANKS];
#pragma unroll
        for (int p = 0; p < DCP_MAX_RANKS; ++p) if (p < W) raw[p] = ld16(slots + ((size_t)p * per + v) * 16);
        x.u = raw[0];
#pragma unroll
        for (int i = 0; i < V; ++i) acc[i] = Tr::up(x.e[i]);
#pragma unroll
        for (int p = 1; p < DCP_MAX_RANKS; ++p) if (p < W) {
          x.u = raw[p];
#pragma unroll
          for (int i = 0; i < V; ++i) acc[i] = red_apply<Acc>(op, acc[i], Tr::up(x.e[i]));
        }
#pragma unroll
        for (int i = 0; i < V; ++i) x.e[i] = Tr::down(red_finish<T>(acc[i], scale, W, op));
        const size_t dst = res_off + (b + v) * 16;
        if (nvls) multimem_st16(c.mc + dst, x.u);
        else {
#pragma unroll
          for (int p = 0; p < DCP_MAX_RANKS; ++p) if (p < W) st16_stream(c.peer[p] + dst, x.u);
        }
      }
    }
    if (!dcp_barrier(c, chan)) return;
    // 4. copy out, slice by slice (chunk index relative to the slice start = the index its owner used)
    char* out = reinterpret_cast<char*>(a.out);
    for (int q = 0; q < W; ++q) {
      const size_t b = per * q < nv ? per * q : nv, e = b + per < nv ? b + per : nv;
      const size_t lo = b * 16, hi = e * 16 < nbytes ? e * 16 : nbytes;
      if (hi > lo) copy_bytes(out + lo, me + res_off + lo, hi - lo, tid, stride, false);
    }
    dcp_gcalls_finish(me, chan);
  }
};

// Rank `root` pushes nbytes to every rank's buffer at `off` (+ parity half when staged), everyone
// copies out.  sym: buffer is the user's symmetric tensor (in place on root, leading barrier needed).
struct BroadcastBody {
  static constexpr int kThreads = kCollThreads;
  struct Args { char* data; size_t off; size_t half_bytes; size_t nbytes; int root; int chan; int sym; };
  static __device__ __forceinline__ void run(const DcpDevCtx& c, const Args& a) {
    char* data = a.data;
    size_t off = a.off;
    size_t half_bytes = a.half_bytes;
    size_t nbytes = a.nbytes;
    int root = a.root;
    int chan = a.chan;
    int sym = a.sym;
  const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (size_t)gridDim.x * blockDim.x;
    if (!sym) off += (size_t)(dcp_gcalls(c.peer[c.rank], chan) & 1u) * half_bytes;
    else if (!dcp_barrier(c, chan)) return;  // peers may still be reading their copy of the tensor
    if (c.rank == root) {
      const char* src = sym ? c.peer[root] + off : data;
      if (c.nvls && c.world > 2 && (off % 16 == 0)) {      // two ranks: one plain peer store per vector runs at link rate, a multicast store does not
        push_all_nvls(c, off, src, nbytes, tid, stride);   // in place on the root for a symmetric tensor: rewrites its own bytes with the same values
      } else {
        for (int p = 0; p < c.world; ++p) {
          if (sym && p == root) continue;
          copy_bytes(c.peer[p] + off, src, nbytes, tid, stride, true);
        }
      }
    }
    if (!dcp_barrier(c, chan)) return;
    if (!sym) {
      if (c.rank != root) copy_bytes(data, c.peer[c.rank] + off, nbytes, tid, stride, false);
      dcp_gcalls_finish(c.peer[c.rank], chan);
    }
  }
};

// Large broadcast with NVLS: a single source cannot fill the switch with multicast stores (measured ~370 GB/s from one GPU),
// while plain peer stores run at link rate (~700 GB/s) and multicast stores from ALL ranks together do as well.  So the root
// SCATTERS: it writes slice p of the message into rank p's buffer with plain stores, sub-chunk by sub-chunk, signalling after
// each one and never waiting; rank p RELAYS every sub-chunk of its slice to all ranks with multimem.st as soon as the root's
// signal for it arrives.  The root's egress carries the message once, every other link direction carries it once, and the two
// phases overlap.  Per-CTA signals: vector v of a slice is handled by the same (CTA, thread) on the root and on the relay.
// Epoch use on (chan, CTA): [leading barrier when sym] + kRelaySub root signals + closing barrier.
constexpr int kRelaySub = 16;
struct BroadcastRelayBody {
  static constexpr int kThreads = kCollThreads;
  struct Args { char* data; size_t off; size_t half_bytes; size_t nbytes; int root; int chan; int sym; };
  static __device__ __forceinline__ void run(const DcpDevCtx& c, const Args& a) {
    char* data = a.data;
    size_t off = a.off;
    const size_t nbytes = a.nbytes;
    const int root = a.root, chan = a.chan, sym = a.sym, W = c.world;
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (size_t)gridDim.x * blockDim.x;
    if (!sym) off += (size_t)(dcp_gcalls(c.peer[c.rank], chan) & 1u) * a.half_bytes;
    else if (!dcp_barrier(c, chan)) return;                       // peers may still be reading their copy of the tensor
    char* me = c.peer[c.rank];
    __syncthreads();                                 // the barrier's epoch update (thread 0) is visible to everyone below
    const uint32_t e0 = *((volatile uint32_t*)dcp_epoch(me, chan, blockIdx.x));
    const size_t nvec = nbytes / 16;                 // the < 16-byte tail travels point to point at the end
    const char* src = (c.rank == root) ? (sym ? me + off : data) : me + off;
    for (int k = 0; k < kRelaySub; ++k) {
      if (c.rank == root) {
        for (int p = 0; p < W; ++p) {
          if (p == root && sym) continue;            // the root's own slice is already in place
          const Span sl = rank_slice(nvec, W, p);
          const size_t len = sl.end - sl.begin, kb = sl.begin + len * k / kRelaySub, ke = sl.begin + len * (k + 1) / kRelaySub;
          size_t v = kb + ((tid + stride - (kb % stride)) % stride);
          for (; v + 3 * stride < ke; v += 4 * stride) {
            const uint4 x0 = ld16(src + v * 16), x1 = ld16(src + (v + stride) * 16), x2 = ld16(src + (v + 2 * stride) * 16), x3 = ld16(src + (v + 3 * stride) * 16);
            char* d = c.peer[p] + off;
            st16_stream(d + v * 16, x0); st16_stream(d + (v + stride) * 16, x1); st16_stream(d + (v + 2 * stride) * 16, x2); st16_stream(d + (v + 3 * stride) * 16, x3);
          }
          for (; v < ke; v += stride) st16_stream(c.peer[p] + off + v * 16, ld16(src + v * 16));
        }
        __syncthreads();
        if ((int)threadIdx.x < W) st_release_sys(dcp_sig(c.peer[threadIdx.x], chan, blockIdx.x, root), e0 + 1 + k);
      }
      // relay this rank's slice (the root relays its own slice too: nobody else has it)
      if (!__syncthreads_and(threadIdx.x != 0 || wait_ge(c, dcp_sig(me, chan, blockIdx.x, root), e0 + 1 + k))) return;   // the root never delivered
      {
        const Span sl = rank_slice(nvec, W, c.rank);
        const size_t len = sl.end - sl.begin, kb = sl.begin + len * k / kRelaySub, ke = sl.begin + len * (k + 1) / kRelaySub;
        const char* from = (c.rank == root) ? src : me + off;
        size_t v = kb + ((tid + stride - (kb % stride)) % stride);
        for (; v + 3 * stride < ke; v += 4 * stride) {
          const uint4 x0 = ld16(from + v * 16), x1 = ld16(from + (v + stride) * 16), x2 = ld16(from + (v + 2 * stride) * 16), x3 = ld16(from + (v + 3 * stride) * 16);
          multimem_st16(c.mc + off + v * 16, x0); multimem_st16(c.mc + off + (v + stride) * 16, x1);
          multimem_st16(c.mc + off + (v + 2 * stride) * 16, x2); multimem_st16(c.mc + off + (v + 3 * stride) * 16, x3);
        }
        for (; v < ke; v += stride) multimem_st16(c.mc + off + v * 16, ld16(from + v * 16));
      }
    }
    if (c.rank == root && (nvec % stride) == tid)
      for (size_t j = nvec * 16; j < nbytes; ++j) { const char x = src[j]; for (int p = 0; p < W; ++p) c.peer[p][off + j] = x; }
    // closing barrier on epoch e0 + kRelaySub + 1 (the epoch word still holds e0: bring it to e0 + kRelaySub first)
    __syncthreads();
    if (threadIdx.x == 0) *((volatile uint32_t*)dcp_epoch(me, chan, blockIdx.x)) = e0 + kRelaySub;
    if (!dcp_barrier(c, chan)) return;
    if (!sym) {
      if (c.rank != root) copy_bytes(data, me + off, nbytes, tid, stride, false);
      dcp_gcalls_finish(me, chan);
    }
  }
};

// Every rank pushes its `nbytes` shard to slot [rank] of every peer's buffer (or only to `root`'s when
// root >= 0: gather).  Staged: copy all W shards out to `out` (contiguous [W][nbytes]) afterwards.
struct AllgatherBody {
  static constexpr int kThreads = kCollThreads;
  struct Args { const char* in; char* out; size_t off; size_t half_bytes; size_t nbytes; size_t slot_stride; size_t out_stride; int root; int chan; int sym; };
  static __device__ __forceinline__ void run(const DcpDevCtx& c, const Args& a) {
    const char* in = a.in;
    char* out = a.out;
    size_t off = a.off;
    size_t half_bytes = a.half_bytes;
    size_t nbytes = a.nbytes;
    size_t slot_stride = a.slot_stride;
    size_t out_stride = a.out_stride;      // bytes between consecutive ranks' shards in `out` (>= nbytes: one chunk of a larger gather)
    int root = a.root;
    int chan = a.chan;
    int sym = a.sym;
  const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (size_t)gridDim.x * blockDim.x;
    if (!sym) off += (size_t)(dcp_gcalls(c.peer[c.rank], chan) & 1u) * half_bytes;
    else if (!dcp_barrier(c, chan)) return;
    const size_t my = off + (size_t)c.rank * slot_stride;
    if (c.nvls && c.world > 2 && root < 0 && (my % 16 == 0)) {   // two ranks: one plain peer store per vector is the faster path
      push_all_nvls(c, my, in, nbytes, tid, stride);        // one store per 16 bytes instead of W: egress drops W-fold
    } else {
      for (int p = 0; p < c.world; ++p) {
        if (root >= 0 && p != root) continue;
        copy_bytes(c.peer[p] + my, in, nbytes, tid, stride, true);
      }
    }
    if (!dcp_barrier(c, chan)) return;
    if (!sym) {
      if (root < 0 || c.rank == root)
        for (int p = 0; p < c.world; ++p) copy_bytes(out + (size_t)p * out_stride, c.peer[c.rank] + off + (size_t)p * slot_stride, nbytes, tid, stride, false);
      dcp_gcalls_finish(c.peer[c.rank], chan);
    }
  }
};

// Rank `root` pushes shard p (in + p*nbytes) to rank p's buffer; everyone copies its shard out.
struct ScatterBody {
  static constexpr int kThreads = kCollThreads;
  struct Args { const char* in; char* out; size_t off; size_t half_bytes; size_t nbytes; size_t in_stride; int root; int chan; };
  static __device__ __forceinline__ void run(const DcpDevCtx& c, const Args& a) {
    const char* in = a.in;
    char* out = a.out;
    size_t off = a.off;
    size_t half_bytes = a.half_bytes;
    size_t nbytes = a.nbytes;
    size_t in_stride = a.in_stride;        // bytes between consecutive ranks' shards in `in` (>= nbytes: one chunk of a larger scatter)
    int root = a.root;
    int chan = a.chan;
  const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (size_t)gridDim.x * blockDim.x;
    off += (size_t)(dcp_gcalls(c.peer[c.rank], chan) & 1u) * half_bytes;
    if (c.rank == root)
      for (int p = 0; p < c.world; ++p) copy_bytes(c.peer[p] + off, in + (size_t)p * in_stride, nbytes, tid, stride, true);
    if (!dcp_barrier(c, chan)) return;
    copy_bytes(out, c.peer[c.rank] + off, nbytes, tid, stride, false);
    dcp_gcalls_finish(c.peer[c.rank], chan);
  }
};

// -------------------------------------------------------------------------------------------------
// reduce_scatter / reduce.
//   root < 0 : reduce_scatter - rank r produces elements [r*n_out, (r+1)*n_out) of the reduction
//   root >= 0: reduce - `root` produces all n_out elements, others only contribute
// sym  (input inside the symmetric heap at offset `off` on every rank): PULL - the owner of a slice reads the W copies of it
//      straight from the peers' tensors (in-switch with NVLS), zero copies; a trailing barrier protects the inputs from reuse.
// !sym (ordinary tensor): PUSH - the staging copy IS the communication: rank r stores piece p of its input into slot [r] of
//      rank p's bounce area (only the owner of a slice ever needs it), one barrier, and every owner reduces its W slots from
//      LOCAL memory.  Slots are padded to 16 bytes, so the reduction is vectorised whatever the alignment of the user tensor;
//      the bounce half comes from the parity of the channel's launch counter (common.cuh), hence no trailing barrier.
//      piece_stride: elements between consecutive pieces of `in` (n_out, or the full shard length when this launch is one chunk
//      of a larger reduce_scatter).
// -------------------------------------------------------------------------------------------------
template <typename T>
struct ReduceScatterBody {
  static constexpr int kThreads = kCollThreads;
  struct Args { const T* in; T* out; size_t off; size_t half_bytes; size_t n_in; size_t n_out; size_t piece_stride; float scale; int op; int root; int chan; int sym; };
  template <bool kLocal>
  static __device__ __forceinline__ void reduce_vec(const DcpDevCtx& c, const char* const* src, size_t v, T* out, float scale, int op, bool out_aligned = true) {
    using Tr = ElemTraits<T>; using Acc = typename Tr::Acc;
    constexpr int V = Vec<T>::N;
    const int W = c.world;
    Acc acc[V]; Vec<T> x;
    uint4 raw[DCP_MAX_RANKS];
#pragma unroll
    for (int p = 0; p < DCP_MAX_RANKS; ++p) if (p < W) raw[p] = kLocal ? ld16(src[p] + v * 16) : ld16_stream(src[p] + v * 16);
    x.u = raw[0];
#pragma unroll
    for (int i = 0; i < V; ++i) acc[i] = Tr::up(x.e[i]);
#pragma unroll
    for (int p = 1; p < DCP_MAX_RANKS; ++p) if (p < W) {
      x.u = raw[p];
#pragma unroll
      for (int i = 0; i < V; ++i) acc[i] = red_apply<Acc>(op, acc[i], Tr::up(x.e[i]));
    }
#pragma unroll
    for (int i = 0; i < V; ++i) x.e[i] = Tr::down(red_finish<T>(acc[i], scale, W, op));
    if (out_aligned) st16(out + v * V, x.u);
    else {
#pragma unroll
      for (int i = 0; i < V; ++i) out[v * V + i] = x.e[i];
    }
  }
  static __device__ __forceinline__ void run(const DcpDevCtx& c, const Args& a) {
    using Tr = ElemTraits<T>; using Acc = typename Tr::Acc;
    constexpr int V = Vec<T>::N;
    const T* in = a.in; T* out = a.out;
    size_t off = a.off;
    const size_t n_out = a.n_out;
    const float scale = a.scale;
    const int op = a.op, root = a.root, chan = a.chan, sym = a.sym, W = c.world;
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (size_t)gridDim.x * blockDim.x;
    char* me = c.peer[c.rank];
    const bool owner = root < 0 || c.rank == root;
    if (!sym) {
      // ---- push: piece p -> slot [my rank] of rank p (reduce: the one piece -> the root)
      off += (size_t)(dcp_gcalls(me, chan) & 1u) * a.half_bytes;
      const size_t slot = (n_out * sizeof(T) + 15) / 16 * 16;
      const size_t pstride = a.piece_stride ? a.piece_stride : n_out;
      for (int p = 0; p < W; ++p) {
        if (root >= 0 && p != root) continue;
        const T* piece = root < 0 ? in + (size_t)p * pstride : in;
        copy_bytes(c.peer[p] + off + (size_t)c.rank * slot, reinterpret_cast<const char*>(piece), n_out * sizeof(T), tid, stride, p != c.rank);
      }
      if (!dcp_barrier(c, chan)) return;
      if (owner) {
        const char* src[DCP_MAX_RANKS];
#pragma unroll
        for (int p = 0; p < DCP_MAX_RANKS; ++p) src[p] = me + off + (size_t)(p < W ? p : 0) * slot;
        // the slots are 16-byte aligned whatever the user tensors look like; vector v (and the tail) is reduced by the (CTA, thread)
        // that copy_bytes made push it on every peer, so the per-CTA barrier above is all the ordering this needs
        const bool al = (reinterpret_cast<uintptr_t>(out) & 15) == 0;
        const size_t nvec = n_out / V;
        for (size_t v = tid; v < nvec; v += stride) reduce_vec<true>(c, src, v, out, scale, op, al);
        if ((nvec % stride) == tid)
          for (size_t i = nvec * V; i < n_out; ++i) {
            Acc acc = Tr::up(reinterpret_cast<const T*>(src[0])[i]);
            for (int p = 1; p < W; ++p) acc = red_apply<Acc>(op, acc, Tr::up(reinterpret_cast<const T*>(me + off + (size_t)p * slot)[i]));
            out[i] = Tr::down(red_finish<T>(acc, scale, W, op));
          }
      }
      dcp_gcalls_finish(me, chan);
      return;
    }
    // ---- pull from the peers' symmetric tensors
    if (!dcp_barrier(c, chan)) return;
    if (owner) {
      const size_t first = root < 0 ? (size_t)c.rank * n_out : 0;
      const size_t soff = off + first * sizeof(T);
      const bool al = ((soff | reinterpret_cast<uintptr_t>(out)) & 15) == 0;
      const size_t nvec = al ? n_out / V : 0;
      bool done_nvls = false;
      if constexpr (NvlsOk<T>::value) {
        if (c.nvls && W > 2 && (op == DCP_SUM || op == DCP_AVG) && nvec > 0) {
          // four in-switch reductions in flight per thread (one is a full round trip through NVSwitch)
          size_t v = tid;
          for (; v + 3 * stride < nvec; v += 4 * stride) {
            uint4 r0 = multimem_ld_reduce_add<T>(c.mc + soff + v * 16), r1 = multimem_ld_reduce_add<T>(c.mc + soff + (v + stride) * 16);
            uint4 r2 = multimem_ld_reduce_add<T>(c.mc + soff + (v + 2 * stride) * 16), r3 = multimem_ld_reduce_add<T>(c.mc + soff + (v + 3 * stride) * 16);
            if (op == DCP_AVG) { r0 = vec_scale<T>(r0, scale); r1 = vec_scale<T>(r1, scale); r2 = vec_scale<T>(r2, scale); r3 = vec_scale<T>(r3, scale); }
            st16(out + v * V, r0); st16(out + (v + stride) * V, r1); st16(out + (v + 2 * stride) * V, r2); st16(out + (v + 3 * stride) * V, r3);
          }
          for (; v < nvec; v += stride) {
            uint4 r = multimem_ld_reduce_add<T>(c.mc + soff + v * 16);
            if (op == DCP_AVG) r = vec_scale<T>(r, scale);
            st16(out + v * V, r);
          }
          done_nvls = true;
        }
      }
      if (!done_nvls) {
        const char* src[DCP_MAX_RANKS];
#pragma unroll
        for (int p = 0; p < DCP_MAX_RANKS; ++p) src[p] = c.peer[p < W ? p : 0] + soff;
        size_t v = tid;
        for (; v + stride < nvec; v += 2 * stride) { reduce_vec<false>(c, src, v, out, scale, op); reduce_vec<false>(c, src, v + stride, out, scale, op); }
        for (; v < nvec; v += stride) reduce_vec<false>(c, src, v, out, scale, op);
      }
      // unaligned slice (n_out * sizeof(T) not a multiple of 16) or tail: element-wise, four independent elements in flight per thread
      size_t i = nvec * V + tid;
      for (; i + 3 * stride < n_out; i += 4 * stride) {
        Acc acc[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) acc[u] = Tr::up(reinterpret_cast<const T*>(c.peer[0] + soff)[i + u * stride]);
        for (int p = 1; p < W; ++p) {
#pragma unroll
          for (int u = 0; u < 4; ++u) acc[u] = red_apply<Acc>(op, acc[u], Tr::up(reinterpret_cast<const T*>(c.peer[p] + soff)[i + u * stride]));
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) out[i + u * stride] = Tr::down(red_finish<T>(acc[u], scale, W, op));
      }
      for (; i < n_out; i += stride) {
        Acc acc = Tr::up(reinterpret_cast<const T*>(c.peer[0] + soff)[i]);
        for (int p = 1; p < W; ++p) acc = red_apply<Acc>(op, acc, Tr::up(reinterpret_cast<const T*>(c.peer[p] + soff)[i]));
        out[i] = Tr::down(red_finish<T>(acc, scale, W, op));
      }
    }
    if (!dcp_barrier(c, chan)) return;                 // the input is the caller's tensor: nobody may still read it after the call
  }
};

// -------------------------------------------------------------------------------------------------
// send / recv through per-source receive slots.  Block b moves its share of the message and owns
// flag word [peer][b]; sequence numbers are kept per (peer, block) in local control memory.
// -------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t* p2p_word(char* heap, size_t region, int peer, int block) {
  return reinterpret_cast<uint32_t*>(heap + region) + (size_t)peer * DCP_P2P_BLOCKS + block;
}

// One launch moves the WHOLE message.  The (source, destination) slot has two halves; each half is cut into DCP_P2P_BLOCKS fixed
// regions of R bytes and region b belongs to block b alone, guarded by block b's own flag / ack words and sequence number.  The
// message is dealt out block-cyclically in pieces of R bytes: piece (ch * grid + b) is block b's chunk `ch`, which travels through
// region b of half (seq & 1).  A half-region is free once the receiver acknowledged the chunk two before, so every block is its own
// two-deep pipeline and no block ever touches memory another block's handshake is responsible for - whatever the sizes and
// grids of earlier messages were.  (One kernel per side instead of one per 8 MiB chunk: that alone took send/recv from about
// half of the link rate to NCCL's.)
__device__ __forceinline__ size_t p2p_region_bytes(const DcpDevCtx& c) { return (c.p2p_slot_bytes / 2 / DCP_P2P_BLOCKS) / 16 * 16; }

struct SendBody {
  static constexpr int kThreads = kCollThreads;
  struct Args { const char* in; size_t nbytes; int dst; };
  static __device__ __forceinline__ void run(const DcpDevCtx& c, const Args& a) {
    const char* in = a.in;
    const size_t nbytes = a.nbytes;
    const int dst = a.dst;
    char* me = c.peer[c.rank];
    const int b = blockIdx.x;
    volatile uint32_t* seqp = p2p_word(me, DcpHeapLayout::kSendSeqOff, dst, b);
    const uint32_t seq0 = *seqp;
    const size_t R = p2p_region_bytes(c), half = c.p2p_slot_bytes / 2, round = R * gridDim.x;
    const size_t chunks = nbytes == 0 ? 1 : (nbytes + round - 1) / round;
    __shared__ int s_ok;
    if (threadIdx.x == 0) s_ok = 1;
    for (size_t ch = 0; ch < chunks; ++ch) {
      const uint32_t seq = seq0 + (uint32_t)ch + 1;
      size_t lo = ch * round + (size_t)b * R; if (lo > nbytes) lo = nbytes;
      size_t hi = lo + R; if (hi > nbytes) hi = nbytes;
      if (threadIdx.x == 0 && !wait_ge(c, p2p_word(me, DcpHeapLayout::kP2PAckOff, dst, b), seq - 2)) s_ok = 0;
      __syncthreads();
      if (!s_ok) return;                    // the receiver never acknowledged: the host sees the error flag
      char* region = c.peer[dst] + c.p2p_slot_off + (size_t)c.rank * c.p2p_slot_bytes + (size_t)(seq & 1) * half + (size_t)b * R;
      copy_bytes(region, in + lo, hi - lo, threadIdx.x, blockDim.x, true);
      __syncthreads();
      if (threadIdx.x == 0) st_release_sys(p2p_word(c.peer[dst], DcpHeapLayout::kP2PFlagOff, c.rank, b), seq);
    }
    if (threadIdx.x == 0) *seqp = seq0 + (uint32_t)chunks;
  }
};

struct RecvBody {
  static constexpr int kThreads = kCollThreads;
  struct Args { char* out; size_t nbytes; int src; };
  static __device__ __forceinline__ void run(const DcpDevCtx& c, const Args& a) {
    char* out = a.out;
    const size_t nbytes = a.nbytes;
    const int src = a.src;
    char* me = c.peer[c.rank];
    const int b = blockIdx.x;
    volatile uint32_t* seqp = p2p_word(me, DcpHeapLayout::kRecvSeqOff, src, b);
    const uint32_t seq0 = *seqp;
    const size_t R = p2p_region_bytes(c), half = c.p2p_slot_bytes / 2, round = R * gridDim.x;
    const size_t chunks = nbytes == 0 ? 1 : (nbytes + round - 1) / round;
    __shared__ int s_ok;
    if (threadIdx.x == 0) s_ok = 1;
    for (size_t ch = 0; ch < chunks; ++ch) {
      const uint32_t seq = seq0 + (uint32_t)ch + 1;
      size_t lo = ch * round + (size_t)b * R; if (lo > nbytes) lo = nbytes;
      size_t hi = lo + R; if (hi > nbytes) hi = nbytes;
      if (threadIdx.x == 0 && !wait_ge(c, p2p_word(me, DcpHeapLayout::kP2PFlagOff, src, b), seq)) s_ok = 0;
      __syncthreads();
      if (!s_ok) return;                    // nothing arrived in time: leave the rest of `out` alone, the host sees the error flag
      const char* region = me + c.p2p_slot_off + (size_t)src * c.p2p_slot_bytes + (size_t)(seq & 1) * half + (size_t)b * R;
      copy_bytes(out + lo, region, hi - lo, threadIdx.x, blockDim.x, false, true);   // the region is rewritten during this kernel: bypass L1
      __syncthreads();
      if (threadIdx.x == 0) st_release_sys(p2p_word(c.peer[src], DcpHeapLayout::kP2PAckOff, c.rank, b), seq);
    }
    if (threadIdx.x == 0) *seqp = seq0 + (uint32_t)chunks;
  }
};

// Both directions of a point-to-point transfer as one body, so that an emulated world can run the sender and the receiver
// of a message in the same launch (mode 0: this rank takes no part).
struct P2PBody {
  static constexpr int kThreads = kCollThreads;
  struct Args { int mode; SendBody::Args s; RecvBody::Args r; };
  static __device__ __forceinline__ void run(const DcpDevCtx& c, const Args& a) {
    if (a.mode == 1) SendBody::run(c, a.s);
    else if (a.mode == 2) RecvBody::run(c, a.r);
  }
};

struct BarrierBody {
  static constexpr int kThreads = 32;
  struct Args { int chan; };
  static __device__ __forceinline__ void run(const DcpDevCtx& c, const Args& a) { dcp_barrier(c, a.chan); }
};

// ---- host side -------------------------------------------------------------------------------------
int grid_for(size_t nbytes, int max_ctas) {
  size_t per_cta = (size_t)kCollThreads * 16 * 4;  // ~4 vectors per thread before adding CTAs
  long g = (long)((nbytes + per_cta - 1) / per_cta);
  if (g < 1) g = 1;
  if (max_ctas <= 0) max_ctas = nbytes >= ((size_t)32 << 20) ? 128 : 64;   // big messages: more CTAs = more loads / stores in flight
  if (g > max_ctas) g = max_ctas;
  if (g > DCP_MAX_BLOCKS) g = DCP_MAX_BLOCKS;
  return (int)g;
}

size_t elem_size(int dtype) {
  switch (dtype) { case DCP_F32: case DCP_I32: return 4; case DCP_BF16: case DCP_F16: return 2; case DCP_I64: case DCP_F64: return 8; case DCP_U8: return 1; }
  return 0;
}

#define DISPATCH_DTYPE(dtype, ...)                                             \
  switch (dtype) {                                                             \
    case DCP_F32: { using T = float; __VA_ARGS__; break; }                     \
    case DCP_BF16: { using T = __nv_bfloat16; __VA_ARGS__; break; }            \
    case DCP_F16: { using T = __half; __VA_ARGS__; break; }                    \
    case DCP_I32: { using T = int32_t; __VA_ARGS__; break; }                   \
    case DCP_I64: { using T = int64_t; __VA_ARGS__; break; }                   \
    case DCP_F64: { using T = double; __VA_ARGS__; break; }                    \
    case DCP_U8: { using T = uint8_t; __VA_ARGS__; break; }                    \
    default: dcp_set_error("unsupported dtype %d", dtype); return -1;          \
  }

}  // namespace

extern "C" {

const DcpDevCtx* dcp_symm_ctx(void* h);
size_t dcp_symm_bytes(void* h);

static inline long heap_offset(const DcpDevCtx& c, const void* p, size_t heap_bytes) {
  const char* base = c.peer[c.rank];
  const char* q = static_cast<const char*>(p);
  if (q >= base && q < base + heap_bytes) return (long)(q - base);
  return -1;
}

// Layout of the staging area.  EVERY signal channel has its own regions: collectives on different channels run on
// different streams at the same time (DDP buckets on the side stream, SyncBatchNorm statistics and user collectives on the
// compute stream), so they must never share a slot.
//   [one-shot slots: channel x 2 halves x DCP_MAX_RANKS x slot]   1/8 of the staging area
//   [bounce, channel 0 (user collectives): 2 halves]              half of the rest
//   [bounce, channels 1..7: 2 halves each]                        the other half, split evenly
static inline size_t oneshot_slot_bytes(const DcpDevCtx& c) {
  size_t s = c.staging_bytes / 8 / DCP_NUM_CHANNELS / (2 * DCP_MAX_RANKS);
  return s / 16 * 16;
}
static inline size_t oneshot_off(const DcpDevCtx& c, int chan) { return c.staging_off + (size_t)chan * oneshot_slot_bytes(c) * 2 * DCP_MAX_RANKS; }
static inline size_t bounce_base(const DcpDevCtx& c) { return c.staging_off + oneshot_slot_bytes(c) * 2 * DCP_MAX_RANKS * DCP_NUM_CHANNELS; }
static inline size_t bounce_total(const DcpDevCtx& c) { return c.staging_bytes - (bounce_base(c) - c.staging_off); }
static inline size_t bounce_half(const DcpDevCtx& c, int chan) {
  const size_t region = chan == 0 ? bounce_total(c) / 2 : bounce_total(c) / 2 / (DCP_NUM_CHANNELS - 1);
  return (region / 2) / 256 * 256;
}
static inline size_t bounce_off(const DcpDevCtx& c, int chan) {
  if (chan == 0) return bounce_base(c);
  const size_t region = (bounce_total(c) / 2 / (DCP_NUM_CHANNELS - 1)) / 256 * 256;
  return bounce_base(c) + (bounce_total(c) / 2) / 256 * 256 + (size_t)(chan - 1) * region;
}

size_t dcp_coll_oneshot_max_bytes(void* h) { return oneshot_slot_bytes(*dcp_symm_ctx(h)); }
size_t dcp_coll_bounce_max_bytes(void* h) { return bounce_half(*dcp_symm_ctx(h), 0); }

static inline bool bad_chan(int chan) {
  if (chan >= 0 && chan < DCP_NUM_CHANNELS) return false;
  dcp_set_error("signal channel %d out of range [0, %d)", chan, DCP_NUM_CHANNELS);
  return true;
}

// algo: 1 one-shot, 2 two-shot P2P, 3 NVLS.  In-place on `buf` (n elements).  If `buf` lies inside the
// symmetric heap the peers read it directly; otherwise it is bounced through the channel's staging area, in as many
// chunks as its size needs.
int dcp_all_reduce(void* h, void* buf, size_t n, int dtype, int op, int algo, int chan, int max_ctas, cudaStream_t stream) {
  const DcpDevCtx& c = *dcp_symm_ctx(h);
  const size_t es = elem_size(dtype), nbytes = n * es;
  if (n == 0) return 0;
  if (bad_chan(chan)) return -1;
  if ((op == DCP_BAND || op == DCP_BOR || op == DCP_BXOR) && (dtype == DCP_F32 || dtype == DCP_BF16 || dtype == DCP_F16 || dtype == DCP_F64)) {
    dcp_set_error("bitwise reduction on a floating dtype"); return -1;
  }
  const float scale = 1.0f / (float)c.world;
  const long off = heap_offset(c, buf, dcp_symm_bytes(h));
  if (algo == 1) {
    if (nbytes > oneshot_slot_bytes(c)) { dcp_set_error("one-shot all_reduce limited to %zu bytes", oneshot_slot_bytes(c)); return -1; }
    int grid = grid_for(nbytes, max_ctas);
    DISPATCH_DTYPE(dtype, (coll_launch<AllreduceOneshotBody<T>>(h, c, {(const T*)buf, (T*)buf, n, scale, op, chan, oneshot_off(c, chan), oneshot_slot_bytes(c)}, grid, stream)));
  } else {
    const bool nvls = algo == 3;
    if (nvls && (!c.nvls || !(dtype == DCP_F32 || dtype == DCP_BF16 || dtype == DCP_F16) || !(op == DCP_SUM || op == DCP_AVG))) {
      dcp_set_error("NVLS all_reduce needs a multicast mapping, f32/bf16/f16 and SUM/AVG"); return -1;
    }
    if (off >= 0 && (off % 16) == 0) {
      int grid = grid_for(nbytes / c.world + 1, max_ctas);
      if (nvls) { DISPATCH_DTYPE(dtype, (coll_launch<AllreduceTwoshotBody<T, true>>(h, c, {nullptr, nullptr, (size_t)off, 0, n, scale, op, chan}, grid, stream))); }
      else { DISPATCH_DTYPE(dtype, (coll_launch<AllreduceTwoshotBody<T, false>>(h, c, {nullptr, nullptr, (size_t)off, 0, n, scale, op, chan}, grid, stream))); }
    } else {
      // ordinary tensor, chunked by the bounce capacity.  With a multicast mapping and more than two ranks: copy into the bounce
      // area, in-switch two-shot there, copy out (measured on 8 GPUs: 928 us at 256 MiB against 1304 us for the push variant,
      // whose W - 1 peer-store streams per rank fall short of the NVLS rate).  Otherwise: the push-based body.
      const size_t half = bounce_half(c, chan);
      if (c.nvls && c.world > 2 && (dtype == DCP_F32 || dtype == DCP_BF16 || dtype == DCP_F16) && (op == DCP_SUM || op == DCP_AVG) && algo != 2) {
        const size_t cap2 = half / es / 16 * 16;
        for (size_t done = 0; done < n; done += cap2) {
          size_t m = n - done < cap2 ? n - done : cap2;
          int grid = grid_for(m * es, max_ctas);
          char* p = (char*)buf + done * es;
          DISPATCH_DTYPE(dtype, (coll_launch<AllreduceTwoshotBody<T, true>>(h, c, {(const T*)p, (T*)p, bounce_off(c, chan), half, m, scale, op, chan}, grid, stream)));
        }
        DCP_CUDA_CHECK(cudaGetLastError());
        return 0;
      }
      const size_t cap_vec = half / 32 - (size_t)c.world - 1;            // W slots + result area = 2 x (vectors + rounding) per chunk
      const size_t cap = cap_vec * 16 / es;
      for (size_t done = 0; done < n; done += cap) {
        size_t m = n - done < cap ? n - done : cap;
        int grid = grid_for(m * es, max_ctas);
        char* p = (char*)buf + done * es;
        DISPATCH_DTYPE(dtype, (coll_launch<AllreduceStagedBody<T>>(h, c, {(const T*)p, (T*)p, bounce_off(c, chan), half, m, scale, op, chan}, grid, stream)));
      }
    }
  }
  DCP_CUDA_CHECK(cudaGetLastError());
  return 0;
}

int dcp_broadcast(void* h, void* buf, size_t nbytes, int root, int chan, int max_ctas, cudaStream_t stream) {
  const DcpDevCtx& c = *dcp_symm_ctx(h);
  if (nbytes == 0) return 0;
  if (bad_chan(chan)) return -1;
  const long off = heap_offset(c, buf, dcp_symm_bytes(h));
  // >= 64 MiB (DCP_BCAST_RELAY_MIN) with a multicast mapping: scatter + relay (BroadcastRelayBody) - measured on 8 GPUs it loses to the
  // root-only multicast push below that (its 16 signalled sub-chunks cost more than they overlap); DCP_BCAST_RELAY=0 turns it off
  static const size_t relay_min = [] { const char* e = getenv("DCP_BCAST_RELAY_MIN"); return e ? (size_t)atol(e) : (size_t)(64u << 20); }();
  static const bool relay_on = [] { const char* e = getenv("DCP_BCAST_RELAY"); return !(e && e[0] == '0'); }();
  if (off >= 0 && off % 16 == 0) {
    if (relay_on && c.nvls && c.world > 2 && nbytes >= relay_min)
      coll_launch<BroadcastRelayBody>(h, c, {(char*)buf, (size_t)off, 0, nbytes, root, chan, 1}, grid_for(nbytes, max_ctas), stream);
    else
      coll_launch<BroadcastBody>(h, c, {(char*)buf, (size_t)off, 0, nbytes, root, chan, 1}, grid_for(nbytes, max_ctas), stream);
  } else {
    const size_t cap = bounce_half(c, chan);
    for (size_t done = 0; done < nbytes; done += cap) {
      size_t m = nbytes - done < cap ? nbytes - done : cap;
      if (relay_on && c.nvls && c.world > 2 && m >= relay_min)
        coll_launch<BroadcastRelayBody>(h, c, {(char*)buf + done, bounce_off(c, chan), cap, m, root, chan, 0}, grid_for(m, max_ctas), stream);
      else
        coll_launch<BroadcastBody>(h, c, {(char*)buf + done, bounce_off(c, chan), cap, m, root, chan, 0}, grid_for(m, max_ctas), stream);
    }
  }
  DCP_CUDA_CHECK(cudaGetLastError());
  return 0;
}

// out: [W][nbytes] contiguous.  root < 0: all_gather; root >= 0: gather to root (out may be null elsewhere).
// A shard larger than the bounce capacity goes through in chunks (shard p's chunk lands at out + p * nbytes + done).
int dcp_all_gather(void* h, const void* in, void* out, size_t nbytes, int root, int chan, int max_ctas, cudaStream_t stream) {
  const DcpDevCtx& c = *dcp_symm_ctx(h);
  if (nbytes == 0) return 0;
  if (bad_chan(chan)) return -1;
  const long off = out ? heap_offset(c, out, dcp_symm_bytes(h)) : -1;
  if (root < 0 && off >= 0 && off % 16 == 0 && nbytes % 16 == 0) {
    coll_launch<AllgatherBody>(h, c, {(const char*)in, nullptr, (size_t)off, 0, nbytes, nbytes, nbytes, -1, chan, 1}, grid_for(nbytes, max_ctas), stream);
  } else {
    const size_t half = bounce_half(c, chan), cap = half / c.world / 256 * 256;
    for (size_t done = 0; done < nbytes; done += cap) {
      size_t m = nbytes - done < cap ? nbytes - done : cap;
      coll_launch<AllgatherBody>(h, c, {(const char*)in + done, out ? (char*)out + done : nullptr, bounce_off(c, chan), half, m, cap, nbytes, root, chan, 0},
                                 grid_for(m, max_ctas), stream);
    }
  }
  DCP_CUDA_CHECK(cudaGetLastError());
  return 0;
}

// in: [W][nbytes] contiguous on root.  Chunked like dcp_all_gather when a shard exceeds the bounce capacity.
int dcp_scatter(void* h, const void* in, void* out, size_t nbytes, int root, int chan, int max_ctas, cudaStream_t stream) {
  const DcpDevCtx& c = *dcp_symm_ctx(h);
  if (nbytes == 0) return 0;
  if (bad_chan(chan)) return -1;
  const size_t half = bounce_half(c, chan);
  for (size_t done = 0; done < nbytes; done += half) {
    size_t m = nbytes - done < half ? nbytes - done : half;
    coll_launch<ScatterBody>(h, c, {in ? (const char*)in + done : nullptr, (char*)out + done, bounce_off(c, chan), half, m, nbytes, root, chan},
                             grid_for(m, max_ctas), stream);
  }
  DCP_CUDA_CHECK(cudaGetLastError());
  return 0;
}

// root < 0: reduce_scatter (in: W*n_out elements, out: n_out).  root >= 0: reduce (in: n_out elements, out on root).
// A non-symmetric input larger than the bounce capacity is reduced in chunks of the output range.
int dcp_reduce_scatter(void* h, const void* in, void* out, size_t n_out, int dtype, int op, int root, int chan, int max_ctas, cudaStream_t stream) {
  const DcpDevCtx& c = *dcp_symm_ctx(h);
  const size_t es = elem_size(dtype);
  if (n_out == 0) return 0;
  if (bad_chan(chan)) return -1;
  if (es == 0) { dcp_set_error("unsupported dtype %d", dtype); return -1; }
  const size_t n_in = root < 0 ? n_out * c.world : n_out;
  const float scale = 1.0f / (float)c.world;
  const long off = heap_offset(c, in, dcp_symm_bytes(h));
  const int sym = off >= 0 && off % 16 == 0;
  if (sym) {
    int grid = grid_for(n_out * es, max_ctas);
    DISPATCH_DTYPE(dtype, (coll_launch<ReduceScatterBody<T>>(h, c, {(const T*)in, (T*)out, (size_t)off, 0, n_in, n_out, 0, scale, op, root, chan, 1}, grid, stream)));
  } else {
    const size_t half = bounce_half(c, chan);
    const size_t cap = (half / c.world - 16) / es / 16 * 16;    // output elements per launch: W padded slots must fit one bounce half
    for (size_t done = 0; done < n_out; done += cap) {
      size_t m = n_out - done < cap ? n_out - done : cap;
      int grid = grid_for(m * es, max_ctas);
      const char* pin = (const char*)in + done * es;
      char* pout = out ? (char*)out + done * es : nullptr;
      DISPATCH_DTYPE(dtype, (coll_launch<ReduceScatterBody<T>>(h, c, {(const T*)pin, (T*)pout, bounce_off(c, chan), half, m * (root < 0 ? c.world : 1), m, n_out, scale, op, root, chan, 0},
                                                               grid, stream)));
    }
  }
  DCP_CUDA_CHECK(cudaGetLastError());
  return 0;
}

// grid of a point-to-point transfer: a function of the message size alone, so that both sides agree
static inline int p2p_grid(const DcpDevCtx& c, size_t nbytes) {
  const size_t R = (c.p2p_slot_bytes / 2 / DCP_P2P_BLOCKS) / 16 * 16;
  long g = (long)((nbytes + R - 1) / R);
  const int cap = c.p2p_blocks > 0 && c.p2p_blocks < DCP_P2P_BLOCKS ? c.p2p_blocks : DCP_P2P_BLOCKS;
  return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

int dcp_send(void* h, const void* in, size_t nbytes, int dst, cudaStream_t stream) {
  const DcpDevCtx& c = *dcp_symm_ctx(h);
  if (dst < 0 || dst >= c.world || dst == c.rank) { dcp_set_error("send: bad destination rank %d", dst); return -1; }
  P2PBody::Args a{}; a.mode = 1; a.s = {(const char*)in, nbytes, dst};
  coll_launch<P2PBody>(h, c, a, p2p_grid(c, nbytes), stream, 1);
  DCP_CUDA_CHECK(cudaGetLastError());
  return 0;
}

int dcp_recv(void* h, void* out, size_t nbytes, int src, cudaStream_t stream) {
  const DcpDevCtx& c = *dcp_symm_ctx(h);
  if (src < 0 || src >= c.world || src == c.rank) { dcp_set_error("recv: bad source rank %d", src); return -1; }
  P2PBody::Args a{}; a.mode = 2; a.r = {(char*)out, nbytes, src};
  coll_launch<P2PBody>(h, c, a, p2p_grid(c, nbytes), stream, 1);
  DCP_CUDA_CHECK(cudaGetLastError());
  return 0;
}

int dcp_barrier_launch(void* h, int chan, cudaStream_t stream) {
  const DcpDevCtx& c = *dcp_symm_ctx(h);
  if (bad_chan(chan)) return -1;
  coll_launch<BarrierBody>(h, c, {chan}, 1, stream);
  DCP_CUDA_CHECK(cudaGetLastError());
  return 0;
}

}  // extern "C"
