// Launch layer of the collective kernels.
//
// Every collective kernel is written as a "body": a struct with an Args POD and a static __device__ run(ctx, args).
// Two __global__ wrappers instantiate it:
//   k_coll<Body>      the production launch - one rank per GPU, ctx and args passed by value;
//   k_coll_emu<Body>  ALL ranks of a small world inside ONE cooperative launch on ONE GPU: blockIdx.y selects the rank,
//                     the per-rank contexts / arguments sit in device arrays.  The body is the same code, so the signal
//                     protocol, the parity double-buffering and the reductions are exercised exactly as on N GPUs.
// The second form exists because kernels that wait on one another must never be issued as SEPARATE launches on one GPU
// (nothing guarantees that they are resident together); a cooperative launch does guarantee it.  It is how the single-GPU
// test tier covers the multi-rank paths (parallel/emulated.py, tests/test_emulated_world_gpu.py).
//
// Host side: coll_launch<Body>(h, ctx, args, grid, stream) launches k_coll<Body> - unless the heap `h` belongs to an
// emulation group (dcp_emu_create), in which case the launch is only RECORDED; dcp_emu_flush() later merges the i-th
// record of every rank into one k_coll_emu launch.
#pragma once
#include <cstring>

#include "coll/coll_common.cuh"

constexpr int kCollThreads = 512;

template <typename Body>
__global__ void __launch_bounds__(Body::kThreads) k_coll(const DcpDevCtx c, const typename Body::Args a) { Body::run(c, a); }

template <typename Body>
__global__ void __launch_bounds__(Body::kThreads) k_coll_emu(const DcpDevCtx* __restrict__ cs, const typename Body::Args* __restrict__ as, int world) {
  if ((int)blockIdx.y < world) Body::run(cs[blockIdx.y], as[blockIdx.y]);
}

constexpr size_t kEmuMaxArgBytes = 192;
struct DcpEmuRecord {
  // launches k_coll_emu<Body> cooperatively over (grid, world) CTAs
  cudaError_t (*launch)(const void* cs_dev, const void* args_dev, int grid, int world, cudaStream_t s);
  size_t arg_bytes;
  int grid;
  int idle_ok;      // ranks that issued nothing for this slot run the body with zeroed args (point-to-point)
  alignas(16) unsigned char args[kEmuMaxArgBytes];
};
// symm.cu: true when heap `h` is a member of an emulation group - the record was queued, nothing was launched.
bool dcp_emu_record(void* h, const DcpEmuRecord& r);

template <typename Body>
cudaError_t dcp_emu_launcher(const void* cs, const void* as, int grid, int world, cudaStream_t s) {
  void* params[] = {(void*)&cs, (void*)&as, (void*)&world};
  return cudaLaunchCooperativeKernel((void*)k_coll_emu<Body>, dim3(grid, world), dim3(Body::kThreads), params, 0, s);
}

template <typename Body>
inline void coll_launch(void* h, const DcpDevCtx& c, const typename Body::Args& a, int grid, cudaStream_t s, int idle_ok = 0) {
  static_assert(sizeof(typename Body::Args) <= kEmuMaxArgBytes, "collective argument block too large for an emulation record");
  if (h != nullptr) {
    DcpEmuRecord r{};
    r.launch = &dcp_emu_launcher<Body>; r.arg_bytes = sizeof(typename Body::Args); r.grid = grid; r.idle_ok = idle_ok;
    memcpy(r.args, &a, sizeof a);
    if (dcp_emu_record(h, r)) return;
  }
  k_coll<Body><<<grid, Body::kThreads, 0, s>>>(c, a);
}
