// Multi-tensor fused optimizers: ONE launch per step for the whole parameter list.
// Adadelta is the reference's optimizer (ref:main.py:124, update rule = torch/optim/adadelta.py:292-302);
// SGD-momentum and Adam(W) are the north-star additions.  Hyper-parameters live in device memory (`hp`)
// so a captured CUDA graph keeps working when StepLR changes the learning rate between epochs.
//
// table: n entries of {param, grad, state1, state2, numel}; blockIdx.y selects the tensor, blockIdx.x
// strides over its elements.  Gradients may be f32 or bf16 (bucket dtype); params/state are f32.
#include "ops/ops_common.cuh"

struct DcpOptEntry { float* p; const void* g; float* s1; float* s2; long long n; };

namespace {
constexpr int kT = 256;
// hp layout: [0] lr [1..5] kind-specific [6] step (float, advanced on device) [7] grad_scale
template <int KIND, typename G>
__device__ __forceinline__ void opt_update(float& p, float g, float& s1, float& s2, const float* hp, float step) {
  const float lr = hp[0];
  if (KIND == 0) {  // Adadelta: hp[1]=rho hp[2]=eps hp[3]=wd
    const float rho = hp[1], eps = hp[2], wd = hp[3];
    if (wd != 0.f) g += wd * p;
    s1 = rho * s1 + (1.f - rho) * g * g;
    float delta = sqrtf(s2 + eps) / sqrtf(s1 + eps) * g;
    s2 = rho * s2 + (1.f - rho) * delta * delta;
    p -= lr * delta;
  } else if (KIND == 1) {  // SGD: hp[1]=momentum hp[2]=dampening hp[3]=wd hp[4]=nesterov
    const float mom = hp[1], damp = hp[2], wd = hp[3];
    if (wd != 0.f) g += wd * p;
    if (mom != 0.f) {
      s1 = step <= 1.f ? g : mom * s1 + (1.f - damp) * g;
      g = hp[4] != 0.f ? g + mom * s1 : s1;
    }
    p -= lr * g;
  } else {  // Adam: hp[1]=b1 hp[2]=b2 hp[3]=eps hp[4]=wd hp[5]=decoupled
    const float b1 = hp[1], b2 = hp[2], eps = hp[3], wd = hp[4];
    if (wd != 0.f) { if (hp[5] != 0.f) p *= 1.f - lr * wd; else g += wd * p; }
    s1 = b1 * s1 + (1.f - b1) * g;
    s2 = b2 * s2 + (1.f - b2) * g * g;
    const float bc1 = 1.f - __powf(b1, step), bc2 = 1.f - __powf(b2, step);
    p -= lr / bc1 * s1 / (sqrtf(s2 / bc2) + eps);
  }
}

template <int KIND, typename G>
__global__ void __launch_bounds__(kT) k_opt(const DcpOptEntry* __restrict__ table, const float* __restrict__ hp) {
  const DcpOptEntry e = table[blockIdx.y];
  const float step = hp[6], gs = hp[7];
  const G* g = static_cast<const G*>(e.g);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < e.n; i += (long long)gridDim.x * blockDim.x) {
    float p = e.p[i], s1 = e.s1 ? e.s1[i] : 0.f, s2 = e.s2 ? e.s2[i] : 0.f;
    opt_update<KIND, G>(p, to_f(g[i]) * gs, s1, s2, hp, step);
    e.p[i] = p;
    if (e.s1) e.s1[i] = s1;
    if (e.s2) e.s2[i] = s2;
  }
}
__global__ void k_opt_advance(float* hp) { hp[6] += 1.f; }
}  // namespace

extern "C" {
// Advance the device step counter (call once per optimizer.step(), before dcp_optim_step).
int dcp_optim_advance(float* hp, cudaStream_t s) {
  k_opt_advance<<<1, 1, 0, s>>>(hp);
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}
// kind: 0 Adadelta, 1 SGD, 2 Adam.  grad_dtype: DCP_F32 or DCP_BF16.  max_numel sizes the grid.
int dcp_optim_step(const void* table, int ntensors, long long max_numel, const float* hp, int kind, int grad_dtype, cudaStream_t s) {
  if (ntensors <= 0) return 0;
  int gx = (int)((max_numel + kT * 4 - 1) / (kT * 4)); if (gx < 1) gx = 1; if (gx > 296) gx = 296;
  dim3 grid(gx, ntensors);
  const DcpOptEntry* t = static_cast<const DcpOptEntry*>(table);
#define LAUNCH(K, G) k_opt<K, G><<<grid, kT, 0, s>>>(t, hp)
  if (grad_dtype == DCP_F32) { if (kind == 0) LAUNCH(0, float); else if (kind == 1) LAUNCH(1, float); else LAUNCH(2, float); }
  else if (grad_dtype == DCP_BF16) { if (kind == 0) LAUNCH(0, __nv_bfloat16); else if (kind == 1) LAUNCH(1, __nv_bfloat16); else LAUNCH(2, __nv_bfloat16); }
  else { dcp_set_error("optimizer gradients must be f32 or bf16"); return -1; }
#undef LAUNCH
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}
}  // extern "C"

// int64 device counter += 1 (the per-step counter that dropout streams hang off inside a CUDA graph)
namespace { __global__ void k_counter_advance(long long* c) { *c += 1; } }
extern "C" int dcp_counter_advance(void* counter, cudaStream_t s) {
  k_counter_advance<<<1, 1, 0, s>>>((long long*)counter);
  DCP_CUDA_CHECK(cudaGetLastError()); return 0;
}
