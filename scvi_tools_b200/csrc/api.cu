// C-ABI housekeeping: error string, ABI version.
#include <stdarg.h>
#include <string.h>

#include "common.cuh"

namespace scvi {
static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

__global__ void __launch_bounds__(1024) colsum_kernel(const float* __restrict__ part, int64_t T, int64_t stride,
                                                     int64_t n, ColsumSegs segs) {
  __shared__ float red[32][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int64_t e = (int64_t)blockIdx.x * 32 + tx;
  float s = 0.f;
  if (e < n) {
#pragma unroll 4
    for (int64_t t = ty; t < T; t += 32) s += part[t * stride + e];
  }
  red[ty][tx] = s;
  __syncthreads();
  if (ty == 0 && e < n) {
    float tot = 0.f;
#pragma unroll
    for (int g = 0; g < 32; ++g) tot += red[g][tx];
    for (int i = 0; i < segs.n; ++i)
      if (segs.dst[i] && e >= segs.begin[i] && e < segs.end[i]) segs.dst[i][e - segs.begin[i]] = tot;
  }
}

// few tiles, many columns: one thread per column
__global__ void colsum_flat_kernel(const float* __restrict__ part, int64_t T, int64_t stride, int64_t n,
                                   ColsumSegs segs) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n) return;
  float tot = 0.f;
  for (int64_t t = 0; t < T; ++t) tot += part[t * stride + e];
  for (int i = 0; i < segs.n; ++i)
    if (segs.dst[i] && e >= segs.begin[i] && e < segs.end[i]) segs.dst[i][e - segs.begin[i]] = tot;
}

int launch_colsum(const float* part, int64_t T, int64_t stride, int64_t n, const ColsumSegs& segs, cudaStream_t st,
                  const char* name) {
  if (n <= 0) return SCVI_OK;
  if (T <= 16 || n >= 65536) {
    colsum_flat_kernel<<<(unsigned)cdiv(n, 256), 256, 0, st>>>(part, T, stride, n, segs);
    SCVI_LAUNCH_CHECK(name);
    return SCVI_OK;
  }
  colsum_kernel<<<(unsigned)cdiv(n, 32), 1024, 0, st>>>(part, T, stride, n, segs);
  SCVI_LAUNCH_CHECK(name);
  return SCVI_OK;
}
}  // namespace scvi

extern "C" {
int scvi_b200_abi_version(void) { return 2; }
const char* scvi_last_error_string(void) { return scvi::g_err; }
}
