// Generic CUDA-core tiled GEMM skeleton with functor operand loaders (implicit-GEMM friendly).
// Used where tensor cores do not pay or the tcgen05 path has no layout for the shape: conv1 (K = 9,
// ref:main.py:32), fc2 (N = 10, ref:main.py:43), the 7x7 stem with C_in = 3, odd sizes.  64x64x16 tiles,
// 256 threads, 4x4 outputs per thread, fp32 accumulate; gridDim.z splits K (atomic epilogue).
#pragma once
#include "ops/ops_common.cuh"

constexpr int SG_BM = 64, SG_BN = 64, SG_BK = 16, SG_T = 256;

// LA(m, k) -> float, LB(k, n) -> float, ST(m, n, acc) stores one output.
template <class LA, class LB, class ST>
__global__ void __launch_bounds__(SG_T) k_simt_gemm(LA la, LB lb, ST st, int M, int N, int K) {
  __shared__ float As[SG_BK][SG_BM + 4], Bs[SG_BK][SG_BN + 4];
  const int m0 = blockIdx.x * SG_BM, n0 = blockIdx.y * SG_BN;
  const int kchunk = (K + gridDim.z - 1) / gridDim.z;
  const int kbeg = blockIdx.z * kchunk, kend = min(K, kbeg + kchunk);
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;  // 16 x 16 threads, each 4 x 4 outputs
  float acc[4][4] = {};
  for (int k0 = kbeg; k0 < kend; k0 += SG_BK) {
    // 64 x 16 elements each for A and B = 1024 -> 4 per thread
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      int e = threadIdx.x + i * SG_T;
      int am = e & 63, ak = e >> 6;     // consecutive threads walk m (A is gathered; no layout assumption)
      int gm = m0 + am, gk = k0 + ak;
      As[ak][am] = (gm < M && gk < kend) ? la(gm, gk) : 0.f;
      int bn = e & 63, bk = e >> 6;
      int gn = n0 + bn; gk = k0 + bk;
      Bs[bk][bn] = (gn < N && gk < kend) ? lb(gk, gn) : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < SG_BK; ++k) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) { a[i] = As[k][ty * 4 + i]; b[i] = Bs[k][tx * 4 + i]; }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int gm = m0 + ty * 4 + i, gn = n0 + tx * 4 + j;
      if (gm < M && gn < N) st(gm, gn, acc[i][j]);
    }
}

template <class LA, class LB, class ST>
inline void launch_simt_gemm(LA la, LB lb, ST st, int M, int N, int K, int ksplit, cudaStream_t s) {
  dim3 grid((M + SG_BM - 1) / SG_BM, (N + SG_BN - 1) / SG_BN, ksplit < 1 ? 1 : ksplit);
  k_simt_gemm<<<grid, SG_T, 0, s>>>(la, lb, st, M, N, K);
}

// choose a K split so that the grid covers the 148 SMs a few times
inline int simt_ksplit(int M, int N, int K) {
  long tiles = (long)((M + SG_BM - 1) / SG_BM) * ((N + SG_BN - 1) / SG_BN);
  long want = (148L * 4 + tiles - 1) / tiles;
  long maxs = (K + 4 * SG_BK - 1) / (4 * SG_BK);
  long sp = want < maxs ? want : maxs;
  return (int)(sp < 1 ? 1 : sp);
}
