// Shared device/host helpers for the SCVI B200 kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <atomic>

#include "scvi_b200.h"

namespace scvi {

void set_error(const char* fmt, ...);
void count_launch(const char* name);  // one call per kernel launch of this library (scvi_launch_count, profiler)

#define SCVI_REQUIRE(cond, ...)          \
  do {                                   \
    if (!(cond)) {                       \
      scvi::set_error(__VA_ARGS__);      \
      return SCVI_EINVAL;                \
    }                                    \
  } while (0)

#define SCVI_LAUNCH_CHECK(name)                                                   \
  do {                                                                            \
    scvi::count_launch(name);                                                     \
    cudaError_t e__ = cudaGetLastError();                                         \
    if (e__ != cudaSuccess) {                                                     \
      scvi::set_error("%s: CUDA launch failed: %s", name, cudaGetErrorString(e__)); \
      return SCVI_ECUDA;                                                          \
    }                                                                             \
  } while (0)

static inline int64_t cdiv(int64_t a, int64_t b) { return (a + b - 1) / b; }

// Opt a kernel into `bytes` of dynamic shared memory, once per (call site, device): `done` is the call site's
// bit mask of devices already configured (the attribute is per device; setting it twice is harmless, so the
// benign race of two host threads on one device needs no lock).
int smem_optin(const void* func, int bytes, std::atomic<uint64_t>& done, const char* who);

// out[e] = sum_t part[t * stride + e], e < n, in a fixed order (deterministic): 32 column lanes x 32
// tile groups per block, serial inside a group, then a fixed-order merge of the 32 group sums.
// The result row is scattered to up to 8 destination segments: dst[i][e - begin[i]] for begin[i] <= e < end[i].
struct ColsumSegs {
  float* dst[8];
  int64_t begin[8], end[8];
  int n;
};
int launch_colsum(const float* part, int64_t T, int64_t stride, int64_t n, const ColsumSegs& segs, cudaStream_t st,
                  const char* name);
static inline ColsumSegs colsum_seg1(float* dst, int64_t n) {
  ColsumSegs s{};
  s.dst[0] = dst; s.begin[0] = 0; s.end[0] = n; s.n = 1;
  return s;
}

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
