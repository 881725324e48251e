// Shared helpers for the sm_100a kernels: status/error plumbing and small device utilities.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/speakingstyle_b200.h"

namespace ss {

void set_error(const char* fmt, ...);
void count_launch(int n = 1);

inline int cuda_fail(cudaError_t e, const char* what) {
  set_error("%s: %s", what, cudaGetErrorString(e));
  return SS_ERR_CUDA;
}

#define SS_REQUIRE(cond, status, ...)  \
  do {                                 \
    if (!(cond)) {                     \
      ::ss::set_error(__VA_ARGS__);    \
      return (status);                 \
    }                                  \
  } while (0)

#define SS_CHECK_LAUNCH(name)                                   \
  do {                                                          \
    cudaError_t e__ = cudaGetLastError();                       \
    if (e__ != cudaSuccess) return ::ss::cuda_fail(e__, name);  \
    ::ss::count_launch();                                       \
  } while (0)

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

__device__ __forceinline__ float apply_act(float v, int act, float slope) {
  switch (act) {
    case SS_ACT_RELU: return fmaxf(v, 0.0f);
    case SS_ACT_TANH: return tanhf(v);
    case SS_ACT_LRELU: return v > 0.0f ? v : v * slope;
    default: return v;
  }
}

// streaming 128-bit accesses (read-once / write-once data should not displace reusable lines in L1)
__device__ __forceinline__ float4 ld_stream(const float4* p) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ void st_stream(float4* p, const float4& v) {
  asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z),
               "f"(v.w)
               : "memory");
}

}  // namespace ss
