// Multi-tensor fused optimizers: ONE launch per step for the whole parameter list.
// Adadelta is the reference's optimizer (ref:main.py:124, update rule = torch/optim/adadelta.py:292-302);
// SGD-momentum and Adam(W) are the north-star additions.  Hyper-parameters live in device memory (`hp`)
// so a captured CUDA graph keeps working when StepLR changes the learning rate between epochs.
//
// table: n entries of {param, grad, state1, state2, numel}; blockIdx.y selects the tensor, blockIdx.x
// strides over its elements.  Gradients may be f32 or bf16 (bucket dtype); params/state are f32.
#include "ops/optim_math.cuh"


namespace {
constexpr int kT = 256;
template <typename G> struct GVec4;
template <> struct GVec4<float> { using type = float4; };
template <> struct GVec4<__nv_bfloat16> { using type = uint2; };
__device__ __forceinline__ void unpack4(const float4& v, float* o) { o[0] = v.x; o[1] = v.y; o[2] = v.z; o[3] = v.w; }
__device__ __forceinline__ void unpack4(const uint2& v, float* o) {
  const __nv_bfloat162 a = *reinterpret_cast<const __nv_bfloat162*>(&v.x), b = *reinterpret_cast<const __nv_bfloat162*>(&v.y);
  o[0] = __low2float(a); o[1] = __high2float(a); o[2] = __low2float(b); o[3] = __high2float(b);
}

template <int KIND, typename G>
__global__ void __launch_bounds__(kT) k_opt(const DcpOptEntry* __restrict__ table, const float* __restrict__ hp) {
  griddep_launch();
  griddep_wait();
  const DcpOptEntry e = table[blockIdx.y];
  const OptHyper H = load_hyper<KIND>(hp);
  const float gs = hp[7];
  const G* g = static_cast<const G*>(e.g);
  const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x, nthr = (long long)gridDim.x * blockDim.x;
  // 16-byte path: four elements per thread and iteration (parameter views inside the flat buckets are 16-byte aligned
  // whenever their offset is a multiple of four elements)
  using GV = typename GVec4<G>::type;
  const bool vec = (((uintptr_t)e.p | (uintptr_t)e.s1 | (uintptr_t)e.s2) & 15) == 0 && ((uintptr_t)e.g & (sizeof(GV) - 1)) == 0;
  const long long n4 = vec ? e.n / 4 : 0;
  for (long long i = tid; i < n4; i += nthr) {
    float4 p4 = reinterpret_cast<float4*>(e.p)[i];
    float4 a4 = e.s1 ? reinterpret_cast<float4*>(e.s1)[i] : make_float4(0.f, 0.f, 0.f, 0.f);
    float4 b4 = e.s2 ? reinterpret_cast<float4*>(e.s2)[i] : make_float4(0.f, 0.f, 0.f, 0.f);
    float gv[4];
    unpack4(reinterpret_cast<const GV*>(g)[i], gv);
    float* pp = reinterpret_cast<float*>(&p4); float* pa = reinterpret_cast<float*>(&a4); float* pb = reinterpret_cast<float*>(&b4);
#pragma unroll
    for (int j = 0; j < 4; ++j) opt_update_h<KIND>(pp[j], gv[j] * gs, pa[j], pb[j], H);
    reinterpret_cast<float4*>(e.p)[i] = p4;
    if (e.s1) reinterpret_cast<float4*>(e.s1)[i] = a4;
    if (e.s2) reinterpret_cast<float4*>(e.s2)[i] = b4;
  }
  for (long long i = n4 * 4 + tid; i < e.n; i += nthr) {
    float p = e.p[i], s1 = e.s1 ? e.s1[i] : 0.f, s2 = e.s2 ? e.s2[i] : 0.f;
    opt_update_h<KIND>(p, to_f(g[i]) * gs, s1, s2, H);
    e.p[i] = p;
    if (e.s1) e.s1[i] = s1;
    if (e.s2) e.s2[i] = s2;
  }
}
__global__ void k_opt_advance(float* hp) { griddep_launch(); griddep_wait(); hp[6] += 1.f; }
}  // namespace

extern "C" {
// Advance the device step counter (call once per optimizer.step(), before dcp_optim_step).
int dcp_optim_advance(float* hp, cudaStream_t s) {
  DCP_CUDA_CHECK(dcp_launch_pdl(k_opt_advance, dim3(1), dim3(1), 0, s, hp)); return 0;
}
// kind: 0 Adadelta, 1 SGD, 2 Adam.  grad_dtype: DCP_F32 or DCP_BF16.  max_numel sizes the grid.
int dcp_optim_step(const void* table, int ntensors, long long max_numel, const float* hp, int kind, int grad_dtype, cudaStream_t s) {
  if (ntensors <= 0) return 0;
  int gx = (int)((max_numel + kT * 4 - 1) / (kT * 4)); if (gx < 1) gx = 1; if (gx > 296) gx = 296;
  dim3 grid(gx, ntensors);
  const DcpOptEntry* t = static_cast<const DcpOptEntry*>(table);
#define LAUNCH(K, G) DCP_CUDA_CHECK(dcp_launch_pdl(k_opt<K, G>, grid, dim3(kT), 0, s, t, hp))
  if (grad_dtype == DCP_F32) { if (kind == 0) LAUNCH(0, float); else if (kind == 1) LAUNCH(1, float); else LAUNCH(2, float); }
  else if (grad_dtype == DCP_BF16) { if (kind == 0) LAUNCH(0, __nv_bfloat16); else if (kind == 1) LAUNCH(1, __nv_bfloat16); else LAUNCH(2, __nv_bfloat16); }
  else { dcp_set_error("optimizer gradients must be f32 or bf16"); return -1; }
#undef LAUNCH
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}
}  // extern "C"

// int64 device counter += 1 (the per-step counter that dropout streams hang off inside a CUDA graph)
namespace { __global__ void k_counter_advance(long long* c) { griddep_launch(); griddep_wait(); *c += 1; } }
extern "C" int dcp_counter_advance(void* counter, cudaStream_t s) {
  DCP_CUDA_CHECK(dcp_launch_pdl(k_counter_advance, dim3(1), dim3(1), 0, s, (long long*)counter)); return 0;
}
