// Optimizer update rules shared by the multi-tensor kernel (ops/optim.cu) and the fused
// reduce-scatter -> update -> all-gather kernel (coll/fused_opt.cu).
// Adadelta is the reference's optimizer (ref:main.py:124; rule = torch/optim/adadelta.py:292-302).
#pragma once
#include "ops/ops_common.cuh"

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

