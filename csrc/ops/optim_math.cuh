// Optimizer update rules shared by the multi-tensor kernel (ops/optim.cu) and the fused
// reduce-scatter -> update -> all-gather kernel (coll/fused_opt.cu).
// Adadelta is the reference's optimizer (ref:main.py:124; rule = torch/optim/adadelta.py:292-302).
#pragma once
#include "ops/ops_common.cuh"

// One row of the device-resident pointer table the multi-tensor optimizer kernels walk (ops/cuda.py packs it as "QQQQq").
struct DcpOptEntry { float* p; const void* g; float* s1; float* s2; long long n; };

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


// Hyper-parameters of one launch, read from the device block ONCE per thread (so nothing of it is re-evaluated per
// element: Adam's bias corrections cost two __powf each).
struct OptHyper {
  float lr, h1, h2, h3, h4, h5, step;
  float inv_bc1, inv_sqrt_bc2;        // Adam: 1 / (1 - b1^t), 1 / sqrt(1 - b2^t)
};
template <int KIND> __device__ __forceinline__ OptHyper load_hyper(const float* __restrict__ hp) {
  OptHyper h;
  h.lr = hp[0]; h.h1 = hp[1]; h.h2 = hp[2]; h.h3 = hp[3]; h.h4 = hp[4]; h.h5 = hp[5]; h.step = hp[6];
  h.inv_bc1 = h.inv_sqrt_bc2 = 1.f;
  if (KIND == 2) { h.inv_bc1 = 1.f / (1.f - __powf(h.h1, h.step)); h.inv_sqrt_bc2 = rsqrtf(1.f - __powf(h.h2, h.step)); }
  return h;
}
// Same update rules as opt_update above, on preloaded hyper-parameters.
template <int KIND>
__device__ __forceinline__ void opt_update_h(float& p, float g, float& s1, float& s2, const OptHyper& h) {
  if (KIND == 0) {  // Adadelta: h1=rho h2=eps h3=wd
    if (h.h3 != 0.f) g += h.h3 * p;
    s1 = h.h1 * s1 + (1.f - h.h1) * g * g;
    const float delta = sqrtf(s2 + h.h2) * rsqrtf(s1 + h.h2) * g;
    s2 = h.h1 * s2 + (1.f - h.h1) * delta * delta;
    p -= h.lr * delta;
  } else if (KIND == 1) {  // SGD: h1=momentum h2=dampening h3=wd h4=nesterov
    if (h.h3 != 0.f) g += h.h3 * p;
    if (h.h1 != 0.f) {
      s1 = h.step <= 1.f ? g : h.h1 * s1 + (1.f - h.h2) * g;
      g = h.h4 != 0.f ? g + h.h1 * s1 : s1;
    }
    p -= h.lr * g;
  } else {  // Adam: h1=b1 h2=b2 h3=eps h4=wd h5=decoupled
    if (h.h4 != 0.f) { if (h.h5 != 0.f) p *= 1.f - h.lr * h.h4; else g += h.h4 * p; }
    s1 = h.h1 * s1 + (1.f - h.h1) * g;
    s2 = h.h2 * s2 + (1.f - h.h2) * g * g;
    p -= h.lr * h.inv_bc1 * s1 / (sqrtf(s2) * h.inv_sqrt_bc2 + h.h3);
  }
}
