// Element-type traits, 128-bit vector helpers and reduction functors for the collective kernels.
#pragma once
#include "common.cuh"

template <typename T> struct ElemTraits;
template <> struct ElemTraits<float> { using Acc = float; static constexpr bool kFloat = true;
  __device__ static Acc up(float v) { return v; } __device__ static float down(Acc a) { return a; } };
template <> struct ElemTraits<double> { using Acc = double; static constexpr bool kFloat = true;
  __device__ static Acc up(double v) { return v; } __device__ static double down(Acc a) { return a; } };
template <> struct ElemTraits<__nv_bfloat16> { using Acc = float; static constexpr bool kFloat = true;
  __device__ static Acc up(__nv_bfloat16 v) { return __bfloat162float(v); } __device__ static __nv_bfloat16 down(Acc a) { return __float2bfloat16_rn(a); } };
template <> struct ElemTraits<__half> { using Acc = float; static constexpr bool kFloat = true;
  __device__ static Acc up(__half v) { return __half2float(v); } __device__ static __half down(Acc a) { return __float2half_rn(a); } };
template <> struct ElemTraits<int32_t> { using Acc = int32_t; static constexpr bool kFloat = false;
  __device__ static Acc up(int32_t v) { return v; } __device__ static int32_t down(Acc a) { return a; } };
template <> struct ElemTraits<int64_t> { using Acc = int64_t; static constexpr bool kFloat = false;
  __device__ static Acc up(int64_t v) { return v; } __device__ static int64_t down(Acc a) { return a; } };
template <> struct ElemTraits<uint8_t> { using Acc = int32_t; static constexpr bool kFloat = false;
  __device__ static Acc up(uint8_t v) { return v; } __device__ static uint8_t down(Acc a) { return (uint8_t)a; } };

// op is a runtime value but warp-uniform; SUM/AVG (the gradient path) is the first, predicted branch.
template <typename A> __device__ __forceinline__ A red_apply(int op, A a, A b) {
  if (op == DCP_SUM || op == DCP_AVG) return a + b;
  if (op == DCP_MAX) return a > b ? a : b;
  if (op == DCP_MIN) return a < b ? a : b;
  if (op == DCP_PROD) return a * b;
  if constexpr (std::is_integral<A>::value) {
    if (op == DCP_BAND) return a & b;
    if (op == DCP_BOR) return a | b;
    if (op == DCP_BXOR) return a ^ b;
  }
  return a;
}

// scale applies only to floating types (AVG); integer AVG divides (floor for non-negative sums).
template <typename T> __device__ __forceinline__ typename ElemTraits<T>::Acc red_finish(typename ElemTraits<T>::Acc a, float scale, int world, int op) {
  if (op != DCP_AVG) return a;
  if constexpr (ElemTraits<T>::kFloat) return a * (typename ElemTraits<T>::Acc)scale;
  else return a / world;
}

template <typename T> struct Vec {
  static constexpr int N = 16 / sizeof(T);
  union { uint4 u; T e[N]; };
};

__device__ __forceinline__ uint4 ld16(const void* p) { return *reinterpret_cast<const uint4*>(p); }
__device__ __forceinline__ void st16(void* p, uint4 v) { *reinterpret_cast<uint4*>(p) = v; }
// streaming variants: peer data is touched once, keep it out of L1
__device__ __forceinline__ uint4 ld16_stream(const void* p) {
  uint4 v;
  asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st16_stream(void* p, uint4 v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// ---- NVLS (multimem) 128-bit reduce-load / broadcast-store -------------------------------------
template <typename T> __device__ __forceinline__ uint4 multimem_ld_reduce_add(const void* mc);
template <> __device__ __forceinline__ uint4 multimem_ld_reduce_add<float>(const void* mc) {
  uint4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(mc) : "memory");
  return v;
}
template <> __device__ __forceinline__ uint4 multimem_ld_reduce_add<__nv_bfloat16>(const void* mc) {
  uint4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(mc) : "memory");
  return v;
}
template <> __device__ __forceinline__ uint4 multimem_ld_reduce_add<__half>(const void* mc) {
  uint4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.f16x2 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(mc) : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st16(void* mc, uint4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

template <typename T> struct NvlsOk { static constexpr bool value = false; };
template <> struct NvlsOk<float> { static constexpr bool value = true; };
template <> struct NvlsOk<__nv_bfloat16> { static constexpr bool value = true; };
template <> struct NvlsOk<__half> { static constexpr bool value = true; };

// scale a 16-byte vector of T in fp32
template <typename T> __device__ __forceinline__ uint4 vec_scale(uint4 raw, float s) {
  if constexpr (ElemTraits<T>::kFloat) {
    Vec<T> v; v.u = raw;
#pragma unroll
    for (int i = 0; i < Vec<T>::N; ++i) v.e[i] = ElemTraits<T>::down(ElemTraits<T>::up(v.e[i]) * (typename ElemTraits<T>::Acc)s);
    return v.u;
  } else {
    return raw;
  }
}
