// Fused BatchNorm(+ReLU), forward and backward, over [M, C] row-major activations - which covers both
// BatchNorm1d on [N, F] (the reference's `self.batchnorm`, ref:main.py:29,40-41, fused with its ReLU)
// and BatchNorm2d on channels-last [N,H,W,C] (M = N*H*W) for the ResNet configs.
//
//   stats kernel : per-channel sum / sum-of-squares, fp32 per thread -> fp64 atomics into a 2*C (or 3*C)
//                  scratch accumulator that the launcher zeroes first
//   apply kernel : mean/rstd from the accumulator, normalise + affine + optional ReLU; CTA 0 also updates
//                  running_mean / running_var (unbiased) / num_batches_tracked and saves mean, rstd
//   backward     : same two-phase shape: reduce sum(dy'), sum(dy' * xhat) -> dx, dgamma, dbeta, with
//                  dy' = dy * (y > 0) when the ReLU is fused
// SyncBatchNorm: the launcher hands the accumulator to the P2P one-shot all_reduce between the two
// phases (parallel/p2p_backend.py), so the statistics cross NVLink as 2*C+1 doubles.
#include "ops/ops_common.cuh"

namespace {

constexpr int kRows = 8;       // row lanes per CTA
constexpr int kCols = 32;      // channels per CTA (one warp wide -> coalesced 128 B rows for f32)

// acc layout: [0,C) sum, [C,2C) sumsq
template <typename T>
__global__ void __launch_bounds__(kRows * kCols) k_bn_stats(const T* __restrict__ x, double* __restrict__ acc, long M, int C, long rows_per_cta) {
  __shared__ float s1[kRows][kCols + 1], s2[kRows][kCols + 1];
  const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
  const int c = blockIdx.x * kCols + cx;
  const long r0 = (long)blockIdx.y * rows_per_cta, r1 = min(M, r0 + rows_per_cta);
  float a = 0.f, b = 0.f;
  if (c < C)
    for (long r = r0 + ry; r < r1; r += kRows) { float v = to_f(x[r * C + c]); a += v; b += v * v; }
  s1[ry][cx] = a; s2[ry][cx] = b;
  __syncthreads();
  if (ry == 0 && c < C) {
#pragma unroll
    for (int i = 1; i < kRows; ++i) { a += s1[i][cx]; b += s2[i][cx]; }
    atomicAdd(&acc[c], (double)a);
    atomicAdd(&acc[C + c], (double)b);
  }
}

// count = number of rows that went into acc (local M, or the global count for SyncBN)
template <typename T>
__global__ void k_bn_apply(const T* __restrict__ x, T* __restrict__ y, const double* __restrict__ acc, const float* __restrict__ gamma,
                           const float* __restrict__ beta, float* __restrict__ run_mean, float* __restrict__ run_var, long long* __restrict__ nbt,
                           float* __restrict__ save_mean, float* __restrict__ save_rstd, long M, int C, double count, float eps, float momentum,
                           int relu, int training) {
  extern __shared__ float sm[];  // scale[C], shift[C]
  float* scale = sm; float* shift = sm + C;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float mean, var;
    if (training) {
      double m = acc[c] / count;
      double v = acc[C + c] / count - m * m;
      if (v < 0) v = 0;
      mean = (float)m; var = (float)v;
    } else { mean = run_mean[c]; var = run_var[c]; }
    float rstd = rsqrtf(var + eps);
    float g = gamma ? gamma[c] : 1.f, bt = beta ? beta[c] : 0.f;
    scale[c] = g * rstd; shift[c] = bt - mean * g * rstd;
    if (blockIdx.x == 0 && training) {
      save_mean[c] = mean; save_rstd[c] = rstd;
      if (run_mean) {
        float unbiased = count > 1 ? var * (float)(count / (count - 1.0)) : var;
        run_mean[c] = (1.f - momentum) * run_mean[c] + momentum * mean;
        run_var[c] = (1.f - momentum) * run_var[c] + momentum * unbiased;
      }
    }
  }
  if (blockIdx.x == 0 && threadIdx.x == 0 && training && nbt) *nbt += 1;
  __syncthreads();
  const long total = M * C;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    int c = i % C;
    float v = to_f(x[i]) * scale[c] + shift[c];
    if (relu) v = fmaxf(v, 0.f);
    y[i] = from_f<T>(v);
  }
}

// acc layout: [0,C) sum(dy'), [C,2C) sum(dy' * xhat)
template <typename T>
__global__ void __launch_bounds__(kRows * kCols) k_bn_bwd_reduce(const T* __restrict__ x, const T* __restrict__ y, const T* __restrict__ dy,
                                                                  const float* __restrict__ mean, const float* __restrict__ rstd,
                                                                  double* __restrict__ acc, long M, int C, long rows_per_cta, int relu) {
  __shared__ float s1[kRows][kCols + 1], s2[kRows][kCols + 1];
  const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
  const int c = blockIdx.x * kCols + cx;
  const long r0 = (long)blockIdx.y * rows_per_cta, r1 = min(M, r0 + rows_per_cta);
  float a = 0.f, b = 0.f;
  if (c < C) {
    const float mu = mean[c], rs = rstd[c];
    for (long r = r0 + ry; r < r1; r += kRows) {
      float g = to_f(dy[r * C + c]);
      if (relu && !(to_f(y[r * C + c]) > 0.f)) g = 0.f;
      a += g; b += g * (to_f(x[r * C + c]) - mu) * rs;
    }
  }
  s1[ry][cx] = a; s2[ry][cx] = b;
  __syncthreads();
  if (ry == 0 && c < C) {
#pragma unroll
    for (int i = 1; i < kRows; ++i) { a += s1[i][cx]; b += s2[i][cx]; }
    atomicAdd(&acc[c], (double)a);
    atomicAdd(&acc[C + c], (double)b);
  }
}

template <typename T>
__global__ void k_bn_bwd_apply(const T* __restrict__ x, const T* __restrict__ y, const T* __restrict__ dy, T* __restrict__ dx,
                               const float* __restrict__ mean, const float* __restrict__ rstd, const float* __restrict__ gamma,
                               const double* __restrict__ acc, float* __restrict__ dgamma, float* __restrict__ dbeta, long M, int C, double count,
                               int relu) {
  extern __shared__ float sm[];  // k1[C] = gamma*rstd, k2[C] = sum_dy/count, k3[C] = sum_dy_xhat/count
  float* k1 = sm; float* k2 = sm + C; float* k3 = sm + 2 * C;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float g = gamma ? gamma[c] : 1.f;
    k1[c] = g * rstd[c];
    k2[c] = (float)(acc[c] / count);
    k3[c] = (float)(acc[C + c] / count);
    if (blockIdx.x == 0) {
      // under SyncBN acc[0..2C) holds the cross-rank sums; parameter gradients use the local copy kept
      // in acc[2C..4C) (DDP averages them afterwards)
      if (dgamma) dgamma[c] = (float)acc[3 * C + c];
      if (dbeta) dbeta[c] = (float)acc[2 * C + c];
    }
  }
  __syncthreads();
  const long total = M * C;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    int c = i % C;
    float g = to_f(dy[i]);
    if (relu && !(to_f(y[i]) > 0.f)) g = 0.f;
    float xhat = (to_f(x[i]) - mean[c]) * rstd[c];
    dx[i] = from_f<T>(k1[c] * (g - k2[c] - xhat * k3[c]));
  }
}

// copies acc[0..2C) into acc[2C..4C) (keeps the local sums for dgamma/dbeta before a cross-rank reduce)
__global__ void k_bn_keep_local(double* acc, int C) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < 2 * C) acc[2 * C + i] = acc[i];
}

// ---- 16-byte variants (C a multiple of the vector width, 16-byte aligned tensors): what the ResNet layers hit ----------
// One thread moves 8 bf16 / 4 f32 channels of a pixel per access.  Reductions: a CTA of 256 threads = (column lanes x row
// lanes), a thread keeps the sums of ITS channels in registers over all rows it visits (4 rows in flight), the row lanes
// meet in shared-memory atomics, one fp64 atomic per channel and CTA goes to the accumulator.
template <typename T> struct BnVec;
template <> struct BnVec<float> {
  static constexpr int N = 4;
  __device__ static void load(const float* p, float* o) { const float4 v = *reinterpret_cast<const float4*>(p); o[0] = v.x; o[1] = v.y; o[2] = v.z; o[3] = v.w; }
  __device__ static void store(float* p, const float* o) { *reinterpret_cast<float4*>(p) = make_float4(o[0], o[1], o[2], o[3]); }
};
template <> struct BnVec<__nv_bfloat16> {
  static constexpr int N = 8;
  __device__ static void load(const __nv_bfloat16* p, float* o) {
    const uint4 v = *reinterpret_cast<const uint4*>(p);
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
    for (int i = 0; i < 4; ++i) { o[2 * i] = __low2float(h[i]); o[2 * i + 1] = __high2float(h[i]); }
  }
  __device__ static void store(__nv_bfloat16* p, const float* o) {
    uint4 v; __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&v);
#pragma unroll
    for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(o[2 * i], o[2 * i + 1]);
    *reinterpret_cast<uint4*>(p) = v;
  }
};
constexpr int kVT = 256;      // threads per CTA of the vector kernels

// kBwd = false: sums of x and x^2.  kBwd = true: sums of dy' and dy' * xhat (dy' = dy masked by y > 0 when relu).
template <typename T, bool kBwd>
__global__ void __launch_bounds__(kVT) k_bn_reduce_vec(const T* __restrict__ x, const T* __restrict__ y, const T* __restrict__ dy,
                                                        const float* __restrict__ mean, const float* __restrict__ rstd, double* __restrict__ acc,
                                                        long M, int C, int cl, long rows_per_cta, int relu) {
  constexpr int VN = BnVec<T>::N;
  __shared__ float sacc[32 * 8 * 2];                     // [column lane][VN][2]
  const int rl = kVT / cl;                              // row lanes
  const int lx = threadIdx.x % cl, ly = threadIdx.x / cl;
  const int vcol = blockIdx.x * cl + lx;                // vector column
  const bool live = vcol * VN < C;
  for (int i = threadIdx.x; i < cl * VN * 2; i += kVT) sacc[i] = 0.f;
  __syncthreads();
  float a[VN], b[VN], mu[VN], rs[VN];
#pragma unroll
  for (int i = 0; i < VN; ++i) { a[i] = 0.f; b[i] = 0.f; mu[i] = 0.f; rs[i] = 1.f; }
  if (kBwd && live) {
#pragma unroll
    for (int i = 0; i < VN; ++i) { mu[i] = mean[vcol * VN + i]; rs[i] = rstd[vcol * VN + i]; }
  }
  const long r0 = (long)blockIdx.y * rows_per_cta, r1 = min(M, r0 + rows_per_cta);
  if (live) {
    for (long r = r0 + ly; r < r1; r += 4L * rl) {
      // branch-free: rows past the end are clamped to the last row and weighted 0, so all 4 (x 3) loads issue back to back
      float xv[4][VN], gv[4][VN], yv[4][VN], wgt[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const long rr = r + (long)u * rl;
        wgt[u] = rr < r1 ? 1.f : 0.f;
        const long o = min(rr, r1 - 1) * C + (long)vcol * VN;
        BnVec<T>::load(x + o, xv[u]);
        if (kBwd) { BnVec<T>::load(dy + o, gv[u]); if (relu) BnVec<T>::load(y + o, yv[u]); }
      }
#pragma unroll
      for (int u = 0; u < 4; ++u)
#pragma unroll
        for (int i = 0; i < VN; ++i) {
          if (kBwd) gv[u][i] *= wgt[u]; else xv[u][i] *= wgt[u];
        }
#pragma unroll
      for (int u = 0; u < 4; ++u)
#pragma unroll
        for (int i = 0; i < VN; ++i) {
          if (!kBwd) { a[i] += xv[u][i]; b[i] += xv[u][i] * xv[u][i]; }
          else {
            float g = gv[u][i];
            if (relu && !(yv[u][i] > 0.f)) g = 0.f;
            a[i] += g; b[i] += g * (xv[u][i] - mu[i]) * rs[i];
          }
        }
    }
  }
  // row lanes that share a warp (cl < 32: lanes lx, lx + cl, ...) fold by shuffle first, then one shared-memory atomic per
  // channel and warp (cl divides 32 or is a multiple of it in every case that matters; otherwise every lane adds for itself)
  const bool fold = cl < 32 && (32 % cl) == 0;
  if (fold) {
#pragma unroll
    for (int i = 0; i < VN; ++i)
      for (int o = cl; o < 32; o <<= 1) { a[i] += __shfl_xor_sync(0xffffffffu, a[i], o); b[i] += __shfl_xor_sync(0xffffffffu, b[i], o); }
  }
  if (live && (!fold || (threadIdx.x & 31) < cl)) {
#pragma unroll
    for (int i = 0; i < VN; ++i) { atomicAdd(&sacc[(lx * VN + i) * 2], a[i]); atomicAdd(&sacc[(lx * VN + i) * 2 + 1], b[i]); }
  }
  __syncthreads();
  (void)rl;
  for (int i = threadIdx.x; i < cl * VN; i += kVT) {
    const int c = blockIdx.x * cl * VN + i;
    if (c < C) { atomicAdd(&acc[c], (double)sacc[i * 2]); atomicAdd(&acc[C + c], (double)sacc[i * 2 + 1]); }
  }
}

// forward apply, 16 bytes per access: same prologue as k_bn_apply (every CTA derives scale / shift from the accumulator,
// CTA 0 updates the running statistics), then y = x * scale[c] + shift[c] (+ReLU)
template <typename T>
__global__ void __launch_bounds__(kVT) k_bn_apply_vec(const T* __restrict__ x, T* __restrict__ y, const double* __restrict__ acc, const float* __restrict__ gamma,
                                                       const float* __restrict__ beta, float* __restrict__ run_mean, float* __restrict__ run_var,
                                                       long long* __restrict__ nbt, float* __restrict__ save_mean, float* __restrict__ save_rstd, long nvec, int C,
                                                       double count, float eps, float momentum, int relu, int training) {
  constexpr int VN = BnVec<T>::N;
  extern __shared__ __align__(16) float sm[];            // scale[C], shift[C]
  float* scale = sm; float* shift = sm + C;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float mean, var;
    if (training) {
      const double m = acc[c] / count;
      double v = acc[C + c] / count - m * m;
      if (v < 0) v = 0;
      mean = (float)m; var = (float)v;
    } else { mean = run_mean[c]; var = run_var[c]; }
    const float rstd = rsqrtf(var + eps), g = gamma ? gamma[c] : 1.f, bt = beta ? beta[c] : 0.f;
    scale[c] = g * rstd; shift[c] = bt - mean * g * rstd;
    if (blockIdx.x == 0 && training) {
      save_mean[c] = mean; save_rstd[c] = rstd;
      if (run_mean) {
        const float unbiased = count > 1 ? var * (float)(count / (count - 1.0)) : var;
        run_mean[c] = (1.f - momentum) * run_mean[c] + momentum * mean;
        run_var[c] = (1.f - momentum) * run_var[c] + momentum * unbiased;
      }
    }
  }
  if (blockIdx.x == 0 && threadIdx.x == 0 && training && nbt) *nbt += 1;
  __syncthreads();
  const int vpr = C / VN;                                // vectors per row
  for (long v = (long)blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += (long)gridDim.x * blockDim.x) {
    const int c0 = (int)((uint32_t)v % (uint32_t)vpr) * VN;       // nvec < 2^32 (checked on the host)
    float t[VN], sc[VN], sh[VN];
    BnVec<T>::load(x + v * VN, t);
#pragma unroll
    for (int i = 0; i < VN; i += 4) {                            // 16-byte shared-memory reads (scalar ones are 8-way bank conflicted)
      *reinterpret_cast<float4*>(&sc[i]) = *reinterpret_cast<const float4*>(&scale[c0 + i]);
      *reinterpret_cast<float4*>(&sh[i]) = *reinterpret_cast<const float4*>(&shift[c0 + i]);
    }
#pragma unroll
    for (int i = 0; i < VN; ++i) { t[i] = t[i] * sc[i] + sh[i]; if (relu) t[i] = fmaxf(t[i], 0.f); }
    BnVec<T>::store(y + v * VN, t);
  }
}
template <typename T>
__global__ void __launch_bounds__(kVT) k_bn_bwd_apply_vec(const T* __restrict__ x, const T* __restrict__ y, const T* __restrict__ dy, T* __restrict__ dx,
                                                           const float* __restrict__ mean, const float* __restrict__ rstd, const float* __restrict__ gamma,
                                                           const double* __restrict__ acc, float* __restrict__ dgamma, float* __restrict__ dbeta, long nvec,
                                                           int C, double count, int relu) {
  constexpr int VN = BnVec<T>::N;
  extern __shared__ __align__(16) float sm[];            // k1 = gamma*rstd, k2 = sum_dy/count, k3 = sum_dy_xhat/count, mean, rstd
  float* k1 = sm; float* k2 = sm + C; float* k3 = sm + 2 * C; float* smu = sm + 3 * C; float* srs = sm + 4 * C;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    const float g = gamma ? gamma[c] : 1.f;
    k1[c] = g * rstd[c]; k2[c] = (float)(acc[c] / count); k3[c] = (float)(acc[C + c] / count); smu[c] = mean[c]; srs[c] = rstd[c];
    if (blockIdx.x == 0) {
      if (dgamma) dgamma[c] = (float)acc[3 * C + c];
      if (dbeta) dbeta[c] = (float)acc[2 * C + c];
    }
  }
  __syncthreads();
  const int vpr = C / VN;
  for (long v = (long)blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += (long)gridDim.x * blockDim.x) {
    const int c0 = (int)((uint32_t)v % (uint32_t)vpr) * VN;
    float xv[VN], gv[VN], yv[VN], q1[VN], q2[VN], q3[VN], qm[VN], qr[VN];
    BnVec<T>::load(x + v * VN, xv);
    BnVec<T>::load(dy + v * VN, gv);
    if (relu) BnVec<T>::load(y + v * VN, yv);
#pragma unroll
    for (int i = 0; i < VN; i += 4) {
      *reinterpret_cast<float4*>(&q1[i]) = *reinterpret_cast<const float4*>(&k1[c0 + i]);
      *reinterpret_cast<float4*>(&q2[i]) = *reinterpret_cast<const float4*>(&k2[c0 + i]);
      *reinterpret_cast<float4*>(&q3[i]) = *reinterpret_cast<const float4*>(&k3[c0 + i]);
      *reinterpret_cast<float4*>(&qm[i]) = *reinterpret_cast<const float4*>(&smu[c0 + i]);
      *reinterpret_cast<float4*>(&qr[i]) = *reinterpret_cast<const float4*>(&srs[c0 + i]);
    }
#pragma unroll
    for (int i = 0; i < VN; ++i) {
      float g = gv[i];
      if (relu && !(yv[i] > 0.f)) g = 0.f;
      const float xhat = (xv[i] - qm[i]) * qr[i];
      xv[i] = q1[i] * (g - q2[i] - xhat * q3[i]);
    }
    BnVec<T>::store(dx + v * VN, xv);
  }
}
template <typename T> bool bn_vec_ok(long M, int C, const void* a, const void* b, const void* c, const void* d) {
  constexpr int VN = BnVec<T>::N;
  auto al = [](const void* p) { return p == nullptr || (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
  return C % VN == 0 && C <= 2048 && M * (long)C >= 65536 && M * (long)C / VN < 4000000000L && al(a) && al(b) && al(c) && al(d);
}
// grid of the vector reductions: column lanes per CTA (<= 32), CTAs over columns, row splits
void reduce_grid_vec(long M, int C, int vn, int* cl, dim3* grid, long* rows_per_cta) {
  const int ncl = C / vn;
  *cl = ncl < 32 ? ncl : 32;
  while (kVT % *cl) --*cl;                              // row lanes = 256 / cl must be whole (ncl is a power of two in practice)
  const int gx = (ncl + *cl - 1) / *cl;
  const int rl = kVT / *cl;
  long want = (148L * 2 + gx - 1) / gx;
  long maxy = (M + rl * 4L - 1) / (rl * 4L);
  long gy = want < maxy ? want : maxy; if (gy < 1) gy = 1; if (gy > 65535) gy = 65535;
  *rows_per_cta = (M + gy - 1) / gy;
  *grid = dim3(gx, (unsigned)gy);
}

void stats_grid(long M, int C, dim3* grid, long* rows_per_cta) {
  int cx = ceil_div(C, kCols);
  long want = (148L * 4 + cx - 1) / cx;           // ~4 CTAs per SM in total
  long maxy = (M + kRows * 4 - 1) / (kRows * 4);  // at least 4 rows per lane
  long gy = want < maxy ? want : maxy; if (gy < 1) gy = 1; if (gy > 65535) gy = 65535;
  *rows_per_cta = (M + gy - 1) / gy;
  *grid = dim3(cx, (unsigned)gy);
}

}  // namespace

extern "C" {

// scratch: >= 4*C doubles.  Phase 1 of forward: zero + accumulate (sum, sumsq).
int dcp_bn_stats(const void* x, void* scratch, long M, int C, int dtype, cudaStream_t s) {
  DCP_CUDA_CHECK(cudaMemsetAsync(scratch, 0, sizeof(double) * 2 * C, s));
  dim3 grid; long rpc;
  bool vec = false;
  DISPATCH_ACT(dtype, (vec = bn_vec_ok<T>(M, C, x, nullptr, nullptr, nullptr)));
  if (vec) {
    int cl; reduce_grid_vec(M, C, dtype == DCP_F32 ? 4 : 8, &cl, &grid, &rpc);
    DISPATCH_ACT(dtype, (k_bn_reduce_vec<T, false><<<grid, kVT, 0, s>>>((const T*)x, nullptr, nullptr, nullptr, nullptr, (double*)scratch, M, C, cl, rpc, 0)));
    DCP_CUDA_CHECK(cudaGetLastError()); return 0;
  }
  stats_grid(M, C, &grid, &rpc);
  DISPATCH_ACT(dtype, (k_bn_stats<T><<<grid, kRows * kCols, 0, s>>>((const T*)x, (double*)scratch, M, C, rpc)));
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}

// Phase 2 of forward.  training=0: uses running stats, scratch may be null.
int dcp_bn_apply(const void* x, void* y, const void* scratch, const float* gamma, const float* beta, float* run_mean, float* run_var, void* nbt,
                 float* save_mean, float* save_rstd, long M, int C, double count, float eps, float momentum, int relu, int training, int dtype,
                 cudaStream_t s) {
  int grid = grid_1d(M * C, 256, 148 * 8);
  size_t smem = sizeof(float) * 2 * C;
  bool vec = false;
  DISPATCH_ACT(dtype, (vec = bn_vec_ok<T>(M, C, x, y, nullptr, nullptr)));
  if (vec) {
    const long nvec = M * C / (dtype == DCP_F32 ? 4 : 8);
    DISPATCH_ACT(dtype, (k_bn_apply_vec<T><<<grid_1d(nvec, kVT, 148 * 8), kVT, smem, s>>>((const T*)x, (T*)y, (const double*)scratch, gamma, beta, run_mean, run_var,
                                                                                         (long long*)nbt, save_mean, save_rstd, nvec, C, count, eps, momentum, relu, training)));
    DCP_CUDA_CHECK(cudaGetLastError()); return 0;
  }
  DISPATCH_ACT(dtype, (k_bn_apply<T><<<grid, 256, smem, s>>>((const T*)x, (T*)y, (const double*)scratch, gamma, beta, run_mean, run_var,
                                                             (long long*)nbt, save_mean, save_rstd, M, C, count, eps, momentum, relu, training)));
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}

// Backward phase 1: zero + accumulate (sum dy', sum dy'*xhat) into scratch[0..2C) and keep a local copy in [2C..4C).
int dcp_bn_bwd_reduce(const void* x, const void* y, const void* dy, const float* mean, const float* rstd, void* scratch, long M, int C, int relu,
                      int dtype, cudaStream_t s) {
  DCP_CUDA_CHECK(cudaMemsetAsync(scratch, 0, sizeof(double) * 2 * C, s));
  dim3 grid; long rpc;
  bool vec = false;
  DISPATCH_ACT(dtype, (vec = bn_vec_ok<T>(M, C, x, y, dy, nullptr)));
  if (vec) {
    int cl; reduce_grid_vec(M, C, dtype == DCP_F32 ? 4 : 8, &cl, &grid, &rpc);
    DISPATCH_ACT(dtype, (k_bn_reduce_vec<T, true><<<grid, kVT, 0, s>>>((const T*)x, (const T*)y, (const T*)dy, mean, rstd, (double*)scratch, M, C, cl, rpc, relu)));
    k_bn_keep_local<<<ceil_div(2 * C, 256), 256, 0, s>>>((double*)scratch, C);
    DCP_CUDA_CHECK(cudaGetLastError()); return 0;
  }
  stats_grid(M, C, &grid, &rpc);
  DISPATCH_ACT(dtype, (k_bn_bwd_reduce<T><<<grid, kRows * kCols, 0, s>>>((const T*)x, (const T*)y, (const T*)dy, mean, rstd, (double*)scratch, M, C, rpc, relu)));
  k_bn_keep_local<<<ceil_div(2 * C, 256), 256, 0, s>>>((double*)scratch, C);
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}

// Backward phase 2.  count: rows behind scratch[0..2C) (global for SyncBN).
int dcp_bn_bwd_apply(const void* x, const void* y, const void* dy, void* dx, const float* mean, const float* rstd, const float* gamma,
                     const void* scratch, float* dgamma, float* dbeta, long M, int C, double count, int relu, int dtype, cudaStream_t s) {
  int grid = grid_1d(M * C, 256, 148 * 8);
  size_t smem = sizeof(float) * 3 * C;
  bool vec = false;
  DISPATCH_ACT(dtype, (vec = bn_vec_ok<T>(M, C, x, y, dy, dx)));
  if (vec) {
    const long nvec = M * C / (dtype == DCP_F32 ? 4 : 8);
    DISPATCH_ACT(dtype, (k_bn_bwd_apply_vec<T><<<grid_1d(nvec, kVT, 148 * 8), kVT, sizeof(float) * 5 * C, s>>>((const T*)x, (const T*)y, (const T*)dy, (T*)dx, mean, rstd,
                                                                                   gamma, (const double*)scratch, dgamma, dbeta, nvec, C, count, relu)));
    DCP_CUDA_CHECK(cudaGetLastError()); return 0;
  }
  DISPATCH_ACT(dtype, (k_bn_bwd_apply<T><<<grid, 256, smem, s>>>((const T*)x, (const T*)y, (const T*)dy, (T*)dx, mean, rstd, gamma,
                                                                 (const double*)scratch, dgamma, dbeta, M, C, count, relu)));
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}

}  // extern "C"

// =====================================================================================================
// Allocation-free BatchNorm forward (training statistics + normalise), for hosts that cannot give us scratch.
//
// Who needs this: dcp_bn_stats / dcp_bn_apply expect the caller to hand in a scratch accumulator (4*C doubles),
// which our Python layer takes from torch's caching allocator.  Some callers have no allocator to ask:
//   * a C/C++ program that links libdcp_cuda.so without PyTorch, so there is no caching allocator behind it;
//   * a process that gave the whole GPU to its own arena at start-up, where a cudaMalloc would fail - and where
//     cudaMalloc / cudaFree would in any case stall the device, which a latency-bound serving or control loop
//     cannot afford in the middle of a step.
// For them the accumulator lives in the library image itself: a file-scope __device__ array of fixed size,
// created once per GPU when the module is loaded there and kept until the process exits
// (DCP_BN_NOALLOC_MAX_C channels -> 2 * 2048 * 8 B = 32 KiB per GPU).
//
// Rules that follow from sharing one array:
//   * C above DCP_BN_NOALLOC_MAX_C is refused (-1, message in dcp_last_error()); nothing is launched.
//   * The array keeps whatever the previous call left in it, so every call first zeroes the 2*C doubles it uses
//     (stream-ordered memset on the caller's stream) before the statistics kernel adds into them.
//   * Two calls that can overlap on the GPU (different streams, or different host threads on different streams)
//     race on the array and corrupt each other's statistics.  Issue them on one stream, or order them with events.
//     Host threads that share ONE stream are safe: the clear and the two launches of a call are enqueued under a lock,
//     so another thread's call cannot slip in between them.
//   * The first call on each device looks up the array's device address (cudaGetSymbolAddress with that device
//     current) and caches it; make that first call before starting any stream capture on that device.
//   * The call itself performs no cudaMalloc / cudaFree and no allocating library call.  With CUDA's default lazy module
//     loading the FIRST call on a device is also when the runtime places this library's device image - the array included -
//     on that GPU; a process that is about to give the whole GPU to its own arena makes one call per device before that
//     (or sets CUDA_MODULE_LOADING=EAGER).
// =====================================================================================================
#include <mutex>

#define DCP_BN_NOALLOC_MAX_C 2048
#define DCP_BN_NOALLOC_MAX_DEVICES 64

// In host code this name is only a handle for the runtime (not a device address): never pass it to a kernel.
static __device__ __align__(256) double g_bn_noalloc_scratch[2 * DCP_BN_NOALLOC_MAX_C];

namespace {
double* g_bn_noalloc_ptr[DCP_BN_NOALLOC_MAX_DEVICES] = {};
std::mutex g_bn_noalloc_mu;
std::mutex g_bn_noalloc_issue_mu;     // held while one call enqueues its clear + two kernels

// Device address of the scratch array on the current device (resolved once per device).
int bn_noalloc_scratch(double** out) {
  int dev = -1;
  DCP_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= DCP_BN_NOALLOC_MAX_DEVICES) { dcp_set_error("bn no-alloc: device index %d out of range", dev); return -1; }
  std::lock_guard<std::mutex> lock(g_bn_noalloc_mu);
  if (g_bn_noalloc_ptr[dev] == nullptr) {
    double* p = nullptr;
    DCP_CUDA_CHECK(cudaGetSymbolAddress((void**)&p, g_bn_noalloc_scratch));
    if (p == nullptr) { dcp_set_error("bn no-alloc: cudaGetSymbolAddress returned a null address"); return -1; }
    g_bn_noalloc_ptr[dev] = p;
  }
  *out = g_bn_noalloc_ptr[dev];
  return 0;
}
}  // namespace

extern "C" {

// Training-mode BatchNorm(+ReLU) forward over x[M, C] (row-major, C innermost) without any caller-provided or
// library-allocated scratch: y = relu?((x - mean) * rstd * gamma + beta); saves mean / rstd for backward and
// updates running_mean / running_var / num_batches_tracked exactly like dcp_bn_stats + dcp_bn_apply do.
// gamma, beta, run_mean, run_var, nbt may be null.  C <= DCP_BN_NOALLOC_MAX_C.  See the block comment above for the
// sharing rules (one call at a time per GPU; first call per device before any stream capture).
// (declared by hand in include/dcp_b200.h - gen_header: skip)
int dcp_bn_forward_noalloc(const void* x, void* y, const float* gamma, const float* beta, float* run_mean, float* run_var, void* nbt,
                           float* save_mean, float* save_rstd, long M, int C, float eps, float momentum, int relu, int dtype, cudaStream_t s) {
  if (C <= 0 || M <= 0) { dcp_set_error("bn no-alloc: empty input (M=%ld, C=%d)", M, C); return -1; }
  if (C > DCP_BN_NOALLOC_MAX_C) {
    dcp_set_error("bn no-alloc: C=%d exceeds the built-in scratch (DCP_BN_NOALLOC_MAX_C=%d); use dcp_bn_stats/dcp_bn_apply with your own scratch",
                  C, DCP_BN_NOALLOC_MAX_C);
    return -1;
  }
  double* scratch = nullptr;
  if (bn_noalloc_scratch(&scratch) != 0) return -1;
  std::lock_guard<std::mutex> issue(g_bn_noalloc_issue_mu);
  // the array still holds the previous call's sums: clear the part this call accumulates into
  DCP_CUDA_CHECK(cudaMemsetAsync(scratch, 0, sizeof(double) * 2 * (size_t)C, s));
  dim3 grid; long rpc; stats_grid(M, C, &grid, &rpc);
  DISPATCH_ACT(dtype, (k_bn_stats<T><<<grid, kRows * kCols, 0, s>>>((const T*)x, scratch, M, C, rpc)));
  DCP_CUDA_CHECK(cudaGetLastError());
  const int agrid = grid_1d(M * C, 256, 148 * 8);
  const size_t smem = sizeof(float) * 2 * C;
  DISPATCH_ACT(dtype, (k_bn_apply<T><<<agrid, 256, smem, s>>>((const T*)x, (T*)y, scratch, gamma, beta, run_mean, run_var, (long long*)nbt,
                                                              save_mean, save_rstd, M, C, (double)M, eps, momentum, relu, 1)));
  DCP_CUDA_CHECK(cudaGetLastError());
  return 0;
}

}  // extern "C"
