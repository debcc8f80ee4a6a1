// On-device minibatch row gather from the HBM-resident count matrix.
// Replaces AnnTorchDataset.__getitem__ + host->device copy
// (reference src/scvi/data/_anntorchdataset.py:115-169).  HBM-bound: 2*4*B*G + 8*B bytes.
#include "common.cuh"

namespace scvi {

// one warp-group of threads per row segment; 16-byte vector copies when alignment allows
template <bool VEC4>
__global__ void gather_rows_kernel(const float* __restrict__ X, int64_t ldx, const int64_t* __restrict__ idx,
                                   int64_t B, int64_t G, float* __restrict__ out, int64_t ldo) {
  // grid.y strides over rows, grid.x * blockDim.x covers the row
  for (int64_t b = blockIdx.y; b < B; b += gridDim.y) {
    const float* src = X + idx[b] * ldx;
    float* dst = out + b * ldo;
    if (VEC4) {
      const int64_t n4 = G >> 2;
      for (int64_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
        float4 v = ld_stream4(src + 4 * i);
        *reinterpret_cast<float4*>(dst + 4 * i) = v;
      }
    } else {
      for (int64_t i = blockIdx.x * blockDim.x + threadIdx.x; i < G; i += (int64_t)gridDim.x * blockDim.x)
        dst[i] = ld_stream(src + i);
    }
  }
}

__global__ void gather_rows_u16_kernel(const uint16_t* __restrict__ X, int64_t ldx,
                                       const int64_t* __restrict__ idx, int64_t B, int64_t G,
                                       float* __restrict__ out, int64_t ldo) {
  for (int64_t b = blockIdx.y; b < B; b += gridDim.y) {
    const uint16_t* src = X + idx[b] * ldx;
    float* dst = out + b * ldo;
    for (int64_t i = blockIdx.x * blockDim.x + threadIdx.x; i < G; i += (int64_t)gridDim.x * blockDim.x)
      dst[i] = (float)src[i];
  }
}

__global__ void gather_i64_kernel(const int64_t* __restrict__ src, const int64_t* __restrict__ idx, int64_t B,
                                  int64_t* __restrict__ out) {
  int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b < B) out[b] = src[idx[b]];
}

}  // namespace scvi

extern "C" {

int scvi_gather_rows(const float* X, int64_t n_rows, int64_t G, int64_t ldx, const int64_t* idx, int64_t B,
                     float* out, int64_t ldo, void* stream) {
  using namespace scvi;
  SCVI_REQUIRE(X && idx && out, "scvi_gather_rows: NULL pointer");
  SCVI_REQUIRE(n_rows >= 0 && G >= 0 && B >= 0 && ldx >= G && ldo >= G, "scvi_gather_rows: bad shape");
  if (B == 0 || G == 0) return SCVI_OK;
  const bool vec = (G % 4 == 0) && (ldx % 4 == 0) && (ldo % 4 == 0) && ((uintptr_t)X % 16 == 0) &&
                   ((uintptr_t)out % 16 == 0);
  const int threads = 256;
  const int64_t per_row = vec ? G / 4 : G;
  // one CTA per row (2 x 16 B in flight per thread at G=2000); very long rows split over grid.x
  dim3 grid((unsigned)std::min<int64_t>(cdiv(per_row, 16 * threads), 64), (unsigned)std::min<int64_t>(B, 65535));
  if (vec)
    gather_rows_kernel<true><<<grid, threads, 0, (cudaStream_t)stream>>>(X, ldx, idx, B, G, out, ldo);
  else
    gather_rows_kernel<false><<<grid, threads, 0, (cudaStream_t)stream>>>(X, ldx, idx, B, G, out, ldo);
  SCVI_LAUNCH_CHECK("scvi_gather_rows");
  return SCVI_OK;
}

int scvi_gather_rows_u16(const uint16_t* X, int64_t n_rows, int64_t G, int64_t ldx, const int64_t* idx,
                         int64_t B, float* out, int64_t ldo, void* stream) {
  using namespace scvi;
  SCVI_REQUIRE(X && idx && out, "scvi_gather_rows_u16: NULL pointer");
  SCVI_REQUIRE(n_rows >= 0 && G >= 0 && B >= 0 && ldx >= G && ldo >= G, "scvi_gather_rows_u16: bad shape");
  if (B == 0 || G == 0) return SCVI_OK;
  dim3 grid((unsigned)std::min<int64_t>(cdiv(G, 16 * 256), 64), (unsigned)std::min<int64_t>(B, 65535));
  gather_rows_u16_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(X, ldx, idx, B, G, out, ldo);
  SCVI_LAUNCH_CHECK("scvi_gather_rows_u16");
  return SCVI_OK;
}

int scvi_gather_i64(const int64_t* src, int64_t n_rows, const int64_t* idx, int64_t B, int64_t* out,
                    void* stream) {
  using namespace scvi;
  SCVI_REQUIRE(src && idx && out, "scvi_gather_i64: NULL pointer");
  SCVI_REQUIRE(n_rows >= 0 && B >= 0, "scvi_gather_i64: bad shape");
  if (B == 0) return SCVI_OK;
  gather_i64_kernel<<<(unsigned)cdiv(B, 256), 256, 0, (cudaStream_t)stream>>>(src, idx, B, out);
  SCVI_LAUNCH_CHECK("scvi_gather_i64");
  return SCVI_OK;
}
}
