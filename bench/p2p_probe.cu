// NVLink peer-copy probe (one process, two GPUs, cudaDeviceEnablePeerAccess): what does it take for ONE source GPU to
// fill its NVLink egress with plain stores, and how do peer loads (pull) compare?  Sweeps grid size, block size and the
// number of independent 16-byte accesses in flight per thread.  The results size the copy loops of csrc/coll/coll.cu.
//
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o /tmp/p2p_probe bench/p2p_probe.cu && /tmp/p2p_probe [MiB]
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s failed: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

__device__ __forceinline__ uint4 ld16(const void* p) { return *reinterpret_cast<const uint4*>(p); }
__device__ __forceinline__ void st16_na(void* p, uint4 v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint4 ld16_na(const void* p) {
  uint4 v;
  asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}

// grid-stride copy, U independent 16-byte accesses per thread and iteration
template <int U, bool kPull>
__global__ void k_copy(char* __restrict__ dst, const char* __restrict__ src, size_t nvec) {
  const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (size_t)gridDim.x * blockDim.x;
  size_t v = tid;
  for (; v + (U - 1) * stride < nvec; v += U * stride) {
    uint4 x[U];
#pragma unroll
    for (int u = 0; u < U; ++u) x[u] = kPull ? ld16_na(src + (v + u * stride) * 16) : ld16(src + (v + u * stride) * 16);
#pragma unroll
    for (int u = 0; u < U; ++u) { if (kPull) *reinterpret_cast<uint4*>(dst + (v + u * stride) * 16) = x[u]; else st16_na(dst + (v + u * stride) * 16, x[u]); }
  }
  for (; v < nvec; v += stride) { uint4 x = ld16(src + v * 16); st16_na(dst + v * 16, x); }
}

// contiguous share per CTA (as k_send): CTA b copies [b * per, (b + 1) * per)
template <int U>
__global__ void k_copy_blocked(char* __restrict__ dst, const char* __restrict__ src, size_t nvec) {
  const size_t per = (nvec + gridDim.x - 1) / gridDim.x, lo = per * blockIdx.x, hi = lo + per < nvec ? lo + per : nvec;
  size_t v = lo + threadIdx.x;
  const size_t stride = blockDim.x;
  for (; v + (U - 1) * stride < hi; v += U * stride) {
    uint4 x[U];
#pragma unroll
    for (int u = 0; u < U; ++u) x[u] = ld16(src + (v + u * stride) * 16);
#pragma unroll
    for (int u = 0; u < U; ++u) st16_na(dst + (v + u * stride) * 16, x[u]);
  }
  for (; v < hi; v += stride) st16_na(dst + v * 16, ld16(src + v * 16));
}

template <typename F> float time_it(F f, int reps) {
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  for (int i = 0; i < 2; ++i) f();
  cudaDeviceSynchronize();
  cudaEventRecord(a);
  for (int i = 0; i < reps; ++i) f();
  cudaEventRecord(b); cudaEventSynchronize(b);
  float ms = 0; cudaEventElapsedTime(&ms, a, b);
  return ms / reps;
}

int main(int argc, char** argv) {
  const size_t mib = argc > 1 ? atol(argv[1]) : 256, bytes = mib << 20, nvec = bytes / 16;
  int ndev = 0; CK(cudaGetDeviceCount(&ndev));
  if (ndev < 2) { printf("needs 2 GPUs\n"); return 0; }
  char *src, *loc, *peer;
  CK(cudaSetDevice(1)); CK(cudaMalloc(&peer, bytes)); CK(cudaMemset(peer, 1, bytes));
  CK(cudaSetDevice(0)); CK(cudaDeviceEnablePeerAccess(1, 0)); CK(cudaMalloc(&src, bytes)); CK(cudaMalloc(&loc, bytes)); CK(cudaMemset(src, 2, bytes));
  printf("{\"probe\": \"cudaMemcpyPeerAsync\", \"MiB\": %zu, \"GBps\": %.1f}\n", mib, bytes / 1e6 / time_it([&] { cudaMemcpyPeerAsync(peer, 1, src, 0, bytes, 0); }, 5));
  const int grids[] = {32, 64, 128, 148, 296, 592};
  const int blocks[] = {256, 512, 1024};
  for (int blk : blocks)
    for (int g : grids) {
#define RUN(NAME, KERN, DST, SRC) { float ms = time_it([&] { KERN<<<g, blk>>>(DST, SRC, nvec); }, 5); \
      printf("{\"probe\": \"%s\", \"grid\": %d, \"block\": %d, \"MiB\": %zu, \"GBps\": %.1f}\n", NAME, g, blk, mib, bytes / 1e6 / ms); }
      RUN("push strided U=1", (k_copy<1, false>), peer, src)
      RUN("push strided U=4", (k_copy<4, false>), peer, src)
      RUN("push strided U=8", (k_copy<8, false>), peer, src)
      RUN("push blocked U=4", (k_copy_blocked<4>), peer, src)
      RUN("push blocked U=8", (k_copy_blocked<8>), peer, src)
      RUN("pull strided U=4", (k_copy<4, true>), loc, peer)
      RUN("pull strided U=8", (k_copy<8, true>), loc, peer)
    }
  CK(cudaGetLastError());
  return 0;
}
