// Shared device/host helpers for the SCVI B200 kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "scvi_b200.h"

namespace scvi {

void set_error(const char* fmt, ...);

#define SCVI_REQUIRE(cond, ...)          \
  do {                                   \
    if (!(cond)) {                       \
      scvi::set_error(__VA_ARGS__);      \
      return SCVI_EINVAL;                \
    }                                    \
  } while (0)

#define SCVI_LAUNCH_CHECK(name)                                                   \
  do {                                                                            \
    cudaError_t e__ = cudaGetLastError();                                         \
    if (e__ != cudaSuccess) {                                                     \
      scvi::set_error("%s: CUDA launch failed: %s", name, cudaGetErrorString(e__)); \
      return SCVI_ECUDA;                                                          \
    }                                                                             \
  } while (0)

static inline int64_t cdiv(int64_t a, int64_t b) { return (a + b - 1) / b; }

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

// streaming 128-bit load that does not pollute L1 (x is read once per kernel)
__device__ __forceinline__ float4 ld_stream4(const float* p) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ float ld_stream(const float* p) {
  float r;
  asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(r) : "l"(p));
  return r;
}

}  // namespace scvi
