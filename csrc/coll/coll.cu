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
    dcp_wait_all(c, chan, e);
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
    dcp_barrier(c, chan);
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
    dcp_barrier(c, chan);
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
__device__ __forceinline__ void copy_bytes(char* dst, const char* src, size_t nbytes, size_t tid, size_t stride, bool stream_store) {
  const uintptr_t mis = reinterpret_cast<uintptr_t>(dst) | reinterpret_cast<uintptr_t>(src);
  const bool al16 = (mis & 15) == 0, al4 = (mis & 3) == 0;
  const size_t nchunk = (nbytes + 15) / 16;
  if (al16) {
    // bulk: 4 independent 16-byte loads in flight per thread (same chunk -> thread mapping as the generic loop)
    const size_t nfull = nbytes / 16;
    size_t c = tid;
    for (; c + 3 * stride < nfull; c += 4 * stride) {
      uint4 x0 = ld16(src + c * 16), x1 = ld16(src + (c + stride) * 16), x2 = ld16(src + (c + 2 * stride) * 16), x3 = ld16(src + (c + 3 * stride) * 16);
      if (stream_store) { st16_stream(dst + c * 16, x0); st16_stream(dst + (c + stride) * 16, x1); st16_stream(dst + (c + 2 * stride) * 16, x2); st16_stream(dst + (c + 3 * stride) * 16, x3); }
      else { st16(dst + c * 16, x0); st16(dst + (c + stride) * 16, x1); st16(dst + (c + 2 * stride) * 16, x2); st16(dst + (c + 3 * stride) * 16, x3); }
    }
    for (; c < nfull; c += stride) { uint4 x = ld16(src + c * 16); if (stream_store) st16_stream(dst + c * 16, x); else st16(dst + c * 16, x); }
    if (nfull < nchunk && (nfull % stride) == tid) for (size_t j = nfull * 16; j < nbytes; ++j) dst[j] = src[j];
    return;
  }
  for (size_t c = tid; c < nchunk; c += stride) {
    const size_t b = c * 16;
    const size_t len = nbytes - b < 16 ? nbytes - b : 16;
    if (al16 && len == 16) {
      uint4 x = ld16(src + b);
      if (stream_store) st16_stream(dst + b, x); else st16(dst + b, x);
    } else if (al4 && (len & 3) == 0) {
      for (size_t j = 0; j < len; j += 4) *reinterpret_cast<uint32_t*>(dst + b + j) = *reinterpret_cast<const uint32_t*>(src + b + j);
    } else {
      for (size_t j = 0; j < len; ++j) dst[b + j] = src[b + j];
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
    else dcp_barrier(c, chan);  // peers may still be reading their copy of the tensor
    if (c.rank == root) {
      const char* src = sym ? c.peer[root] + off : data;
      if (c.nvls && (off % 16 == 0)) {
        push_all_nvls(c, off, src, nbytes, tid, stride);   // in place on the root for a symmetric tensor: rewrites its own bytes with the same values
      } else {
        for (int p = 0; p < c.world; ++p) {
          if (sym && p == root) continue;
          copy_bytes(c.peer[p] + off, src, nbytes, tid, stride, true);
        }
      }
    }
    dcp_barrier(c, chan);
    if (!sym) {
      if (c.rank != root) copy_bytes(data, c.peer[c.rank] + off, nbytes, tid, stride, false);
      dcp_gcalls_finish(c.peer[c.rank], chan);
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
    else dcp_barrier(c, chan);
    const size_t my = off + (size_t)c.rank * slot_stride;
    if (c.nvls && root < 0 && (my % 16 == 0)) {
      push_all_nvls(c, my, in, nbytes, tid, stride);        // one store per 16 bytes instead of W: egress drops W-fold
    } else {
      for (int p = 0; p < c.world; ++p) {
        if (root >= 0 && p != root) continue;
        copy_bytes(c.peer[p] + my, in, nbytes, tid, stride, true);
      }
    }
    dcp_barrier(c, chan);
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
    dcp_barrier(c, chan);
    copy_bytes(out, c.peer[c.rank] + off, nbytes, tid, stride, false);
    dcp_gcalls_finish(c.peer[c.rank], chan);
  }
};

// -------------------------------------------------------------------------------------------------
// reduce_scatter / reduce: stage own full input, then pull-reduce the slice(s) this rank owns.
//   root < 0 : reduce_scatter - rank r produces elements [r*n_out, (r+1)*n_out) of the reduction
//   root >= 0: reduce - `root` produces all n_in elements, others only contribute
// sym: input already at heap offset `off` on every rank (trailing barrier protects it from reuse).
// -------------------------------------------------------------------------------------------------
template <typename T>
struct ReduceScatterBody {
  static constexpr int kThreads = kCollThreads;
  struct Args { const T* in; T* out; size_t off; size_t half_bytes; size_t n_in; size_t n_out; size_t piece_stride; float scale; int op; int root; int chan; int sym; };
  static __device__ __forceinline__ void run(const DcpDevCtx& c, const Args& a) {
    const T* in = a.in;
    T* out = a.out;
    size_t off = a.off;
    size_t half_bytes = a.half_bytes;
    size_t n_in = a.n_in;
    size_t n_out = a.n_out;
    float scale = a.scale;
    int op = a.op;
    int root = a.root;
    int chan = a.chan;
    int sym = a.sym;
  using Tr = ElemTraits<T>; using Acc = typename Tr::Acc;
    constexpr int V = Vec<T>::N;
    const int W = c.world;
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (size_t)gridDim.x * blockDim.x;
    // staged input: the bounce half comes from the parity of the channel's LAUNCH counter (the reduce partition differs from
    // the copy-in partition, so the per-CTA call counters of the other staged kernels would not do).  A rank can overwrite a
    // half only two launches later, and by then every peer has passed a barrier of the launch in between - i.e. has completed,
    // in stream order, the launch that read this half.  Hence no trailing barrier on the staged path.
    // piece_stride != 0: the input is W pieces of n_out elements, piece_stride elements apart (one chunk of a larger
    // reduce_scatter); they are packed back to back in the bounce buffer.
    if (!sym) {
      off += (size_t)(dcp_gcalls(c.peer[c.rank], chan) & 1u) * half_bytes;
      const size_t a_piece_stride = a.piece_stride;
      if (a_piece_stride == 0) copy_bytes(c.peer[c.rank] + off, reinterpret_cast<const char*>(in), n_in * sizeof(T), tid, stride, false);
      else for (int p = 0; p < W; ++p)
        copy_bytes(c.peer[c.rank] + off + (size_t)p * n_out * sizeof(T), reinterpret_cast<const char*>(in + (size_t)p * a_piece_stride), n_out * sizeof(T), tid, stride, false);
    }
    dcp_barrier(c, chan);
    if (root < 0 || c.rank == root) {
      const size_t first = root < 0 ? (size_t)c.rank * n_out : 0;
      const size_t soff = off + first * sizeof(T);
      const bool al = ((soff | reinterpret_cast<uintptr_t>(out)) & 15) == 0;
      const size_t nvec = al ? n_out / V : 0;
      bool done_nvls = false;
      if constexpr (NvlsOk<T>::value) {
        if (c.nvls && (op == DCP_SUM || op == DCP_AVG) && nvec > 0) {
          for (size_t v = tid; v < nvec; v += stride) {
            uint4 r = multimem_ld_reduce_add<T>(c.mc + soff + v * 16);      // in-switch reduction of the W copies
            if (op == DCP_AVG) r = vec_scale<T>(r, scale);
            st16(out + v * V, r);
          }
          done_nvls = true;
        }
      }
      for (size_t v = done_nvls ? nvec : tid; v < nvec; v += stride) {
        Acc acc[V]; Vec<T> x;
        uint4 raw[DCP_MAX_RANKS];
#pragma unroll
        for (int p = 0; p < DCP_MAX_RANKS; ++p) if (p < W) raw[p] = ld16_stream(c.peer[p] + soff + v * 16);
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
        st16(out + v * V, x.u);
      }
      for (size_t i = nvec * V + tid; i < n_out; i += stride) {
        Acc acc = Tr::up(reinterpret_cast<const T*>(c.peer[0] + soff)[i]);
        for (int p = 1; p < W; ++p) acc = red_apply<Acc>(op, acc, Tr::up(reinterpret_cast<const T*>(c.peer[p] + soff)[i]));
        out[i] = Tr::down(red_finish<T>(acc, scale, W, op));
      }
    }
    if (sym) dcp_barrier(c, chan);                 // the input is the caller's tensor: nobody may still read it after the call
    else dcp_gcalls_finish(c.peer[c.rank], chan);
  }
};

// -------------------------------------------------------------------------------------------------
// send / recv through per-source receive slots.  Block b moves its share of the message and owns
// flag word [peer][b]; sequence numbers are kept per (peer, block) in local control memory.
// -------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t* p2p_word(char* heap, size_t region, int peer, int block) {
  return reinterpret_cast<uint32_t*>(heap + region) + (size_t)peer * DCP_P2P_BLOCKS + block;
}

struct SendBody {
  static constexpr int kThreads = kCollThreads;
  struct Args { const char* in; size_t nbytes; int dst; };
  static __device__ __forceinline__ void run(const DcpDevCtx& c, const Args& a) {
    const char* in = a.in;
    size_t nbytes = a.nbytes;
    int dst = a.dst;
  char* me = c.peer[c.rank];
    const int b = blockIdx.x;
    volatile uint32_t* seqp = p2p_word(me, DcpHeapLayout::kSendSeqOff, dst, b);
    const uint32_t seq = *seqp + 1;
    // two half-slots per (source, destination): message seq uses half (seq & 1), which is free once the receiver
    // acknowledged message seq - 2 - so a stream of chunks pipelines instead of ping-ponging on acks
    if (threadIdx.x == 0) wait_ge(c, p2p_word(me, DcpHeapLayout::kP2PAckOff, dst, b), seq - 2);
    __syncthreads();
    char* slot = c.peer[dst] + c.p2p_slot_off + (size_t)c.rank * c.p2p_slot_bytes + (size_t)(seq & 1) * (c.p2p_slot_bytes / 2);
    // contiguous share per block so that flags cover disjoint ranges
    size_t per = ((nbytes + gridDim.x - 1) / gridDim.x + 15) / 16 * 16;
    size_t lo = per * b; if (lo > nbytes) lo = nbytes;
    size_t hi = lo + per; if (hi > nbytes) hi = nbytes;
    copy_bytes(slot + lo, in + lo, hi - lo, threadIdx.x, blockDim.x, true);
    __syncthreads();
    if (threadIdx.x == 0) {
      st_release_sys(p2p_word(c.peer[dst], DcpHeapLayout::kP2PFlagOff, c.rank, b), seq);
      *seqp = seq;
    }
  }
};

struct RecvBody {
  static constexpr int kThreads = kCollThreads;
  struct Args { char* out; size_t nbytes; int src; };
  static __device__ __forceinline__ void run(const DcpDevCtx& c, const Args& a) {
    char* out = a.out;
    size_t nbytes = a.nbytes;
    int src = a.src;
  char* me = c.peer[c.rank];
    const int b = blockIdx.x;
    volatile uint32_t* seqp = p2p_word(me, DcpHeapLayout::kRecvSeqOff, src, b);
    const uint32_t seq = *seqp + 1;
    if (threadIdx.x == 0) wait_ge(c, p2p_word(me, DcpHeapLayout::kP2PFlagOff, src, b), seq);
    __syncthreads();
    const char* slot = me + c.p2p_slot_off + (size_t)src * c.p2p_slot_bytes + (size_t)(seq & 1) * (c.p2p_slot_bytes / 2);
    size_t per = ((nbytes + gridDim.x - 1) / gridDim.x + 15) / 16 * 16;
    size_t lo = per * b; if (lo > nbytes) lo = nbytes;
    size_t hi = lo + per; if (hi > nbytes) hi = nbytes;
    copy_bytes(out + lo, slot + lo, hi - lo, threadIdx.x, blockDim.x, false);
    __syncthreads();
    if (threadIdx.x == 0) {
      st_release_sys(p2p_word(c.peer[src], DcpHeapLayout::kP2PAckOff, c.rank, b), seq);
      *seqp = seq;
    }
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
      const size_t half = bounce_half(c, chan), cap = half / es / 16 * 16;
      for (size_t done = 0; done < n; done += cap) {
        size_t m = n - done < cap ? n - done : cap;
        int grid = grid_for(m * es, max_ctas);
        char* p = (char*)buf + done * es;
        if (nvls) { DISPATCH_DTYPE(dtype, (coll_launch<AllreduceTwoshotBody<T, true>>(h, c, {(const T*)p, (T*)p, bounce_off(c, chan), half, m, scale, op, chan}, grid, stream))); }
        else { DISPATCH_DTYPE(dtype, (coll_launch<AllreduceTwoshotBody<T, false>>(h, c, {(const T*)p, (T*)p, bounce_off(c, chan), half, m, scale, op, chan}, grid, stream))); }
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
  if (off >= 0 && off % 16 == 0) {
    coll_launch<BroadcastBody>(h, c, {(char*)buf, (size_t)off, 0, nbytes, root, chan, 1}, grid_for(nbytes, max_ctas), stream);
  } else {
    const size_t cap = bounce_half(c, chan);
    for (size_t done = 0; done < nbytes; done += cap) {
      size_t m = nbytes - done < cap ? nbytes - done : cap;
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
    const size_t pieces = root < 0 ? (size_t)c.world : 1;
    const size_t cap = half / pieces / es / 16 * 16;          // output elements per launch (keeps every piece 16-byte aligned in the buffer)
    for (size_t done = 0; done < n_out; done += cap) {
      size_t m = n_out - done < cap ? n_out - done : cap;
      int grid = grid_for(m * es, max_ctas);
      const char* pin = (const char*)in + done * es;
      char* pout = out ? (char*)out + done * es : nullptr;
      const size_t stride = (root < 0 && m != n_out) ? n_out : 0;   // one chunk of a larger reduce_scatter: W pieces, n_out elements apart
      DISPATCH_DTYPE(dtype, (coll_launch<ReduceScatterBody<T>>(h, c, {(const T*)pin, (T*)pout, bounce_off(c, chan), half, m * pieces, m, stride, scale, op, root, chan, 0},
                                                               grid, stream)));
    }
  }
  DCP_CUDA_CHECK(cudaGetLastError());
  return 0;
}

int dcp_send(void* h, const void* in, size_t nbytes, int dst, cudaStream_t stream) {
  const DcpDevCtx& c = *dcp_symm_ctx(h);
  const size_t half = c.p2p_slot_bytes / 2;
  for (size_t done = 0; done < nbytes || (nbytes == 0 && done == 0); done += half) {
    size_t m = nbytes - done < half ? nbytes - done : half;
    P2PBody::Args a{}; a.mode = 1; a.s = {(const char*)in + done, m, dst};
    coll_launch<P2PBody>(h, c, a, DCP_P2P_BLOCKS, stream, 1);
    if (nbytes == 0) break;
  }
  DCP_CUDA_CHECK(cudaGetLastError());
  return 0;
}

int dcp_recv(void* h, void* out, size_t nbytes, int src, cudaStream_t stream) {
  const DcpDevCtx& c = *dcp_symm_ctx(h);
  const size_t half = c.p2p_slot_bytes / 2;
  for (size_t done = 0; done < nbytes || (nbytes == 0 && done == 0); done += half) {
    size_t m = nbytes - done < half ? nbytes - done : half;
    P2PBody::Args a{}; a.mode = 2; a.r = {(char*)out + done, m, src};
    coll_launch<P2PBody>(h, c, a, DCP_P2P_BLOCKS, stream, 1);
    if (nbytes == 0) break;
  }
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
