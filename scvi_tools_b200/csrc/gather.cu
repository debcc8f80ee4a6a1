// On-device minibatch row gather from the HBM-resident count matrix.
// Replaces AnnTorchDataset.__getitem__ + host->device copy
// (reference src/scvi/data/_anntorchdataset.py:115-169).  HBM-bound: 2*4*B*G + 8*B bytes.
#include "common.cuh"

namespace scvi {

// GR_ROWS rows per CTA pass: the row ids are read first, then every thread issues the 16-byte loads of its column
// piece for all of them before the first store -- 64 bytes in flight per thread and iteration instead of 16 (measured
// alone, 8192 random rows of 2000 floats: 37.3 us with one row per pass, 36.7 with two, 34.9 with four).
constexpr int GR_ROWS = 4;
template <bool VEC4>
__global__ void gather_rows_kernel(const float* __restrict__ X, int64_t ldx, const int64_t* __restrict__ idx,
                                   int64_t B, int64_t G, float* __restrict__ out, int64_t ldo) {
  // grid.y strides over groups of GR_ROWS rows, grid.x * blockDim.x covers the row
  for (int64_t b0 = (int64_t)blockIdx.y * GR_ROWS; b0 < B; b0 += (int64_t)gridDim.y * GR_ROWS) {
    const float* src[GR_ROWS];
#pragma unroll
    for (int j = 0; j < GR_ROWS; ++j) src[j] = X + idx[min(b0 + j, B - 1)] * ldx;  // past the end: a duplicate read
    if (VEC4) {
      const int64_t n4 = G >> 2;
      for (int64_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
        float4 v[GR_ROWS];
#pragma unroll
        for (int j = 0; j < GR_ROWS; ++j) v[j] = ld_stream4(src[j] + 4 * i);
#pragma unroll
        for (int j = 0; j < GR_ROWS; ++j)
          if (b0 + j < B) *reinterpret_cast<float4*>(out + (b0 + j) * ldo + 4 * i) = v[j];
      }
    } else {
      for (int64_t i = blockIdx.x * blockDim.x + threadIdx.x; i < G; i += (int64_t)gridDim.x * blockDim.x) {
        float v[GR_ROWS];
#pragma unroll
        for (int j = 0; j < GR_ROWS; ++j) v[j] = ld_stream(src[j] + i);
#pragma unroll
        for (int j = 0; j < GR_ROWS; ++j)
          if (b0 + j < B) out[(b0 + j) * ldo + i] = v[j];
      }
    }
  }
}

// Compact resident / staged counts: T = uint8_t or uint16_t (the narrowest integer type that holds the
// counts losslessly).  out[b, :] = (float)X[idx ? idx[b] : b, :] -- exact, every such integer is an fp32 value.
// VEC: one 16-byte load (16 / 8 counts) per thread and 4 / 2 float4 stores; rows must be 16-byte aligned.
template <typename T, bool VEC>
__global__ void gather_rows_int_kernel(const T* __restrict__ X, int64_t ldx, const int64_t* __restrict__ idx,
                                       int64_t B, int64_t G, float* __restrict__ out, int64_t ldo) {
  constexpr int PER = 16 / (int)sizeof(T);  // counts per 16-byte load
  for (int64_t b = blockIdx.y; b < B; b += gridDim.y) {
    const T* src = X + (idx ? idx[b] : b) * ldx;
    float* dst = out + b * ldo;
    if (VEC) {
      const int64_t nv = G / PER;
      for (int64_t i = blockIdx.x * blockDim.x + threadIdx.x; i < nv; i += (int64_t)gridDim.x * blockDim.x) {
        uint4 raw;
        asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                     : "=r"(raw.x), "=r"(raw.y), "=r"(raw.z), "=r"(raw.w)
                     : "l"(src + PER * i));
        const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
        float4* d4 = reinterpret_cast<float4*>(dst + PER * i);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          if (sizeof(T) == 1) {
            d4[k] = make_float4((float)(w[k] & 0xffu), (float)((w[k] >> 8) & 0xffu), (float)((w[k] >> 16) & 0xffu),
                                (float)(w[k] >> 24));
          } else if ((k & 1) == 0) {
            d4[k >> 1] = make_float4((float)(w[k] & 0xffffu), (float)(w[k] >> 16), (float)(w[k + 1] & 0xffffu),
                                     (float)(w[k + 1] >> 16));
          }
        }
      }
      for (int64_t i = nv * PER + blockIdx.x * blockDim.x + threadIdx.x; i < G; i += (int64_t)gridDim.x * blockDim.x)
        dst[i] = (float)src[i];
    } else {
      for (int64_t i = blockIdx.x * blockDim.x + threadIdx.x; i < G; i += (int64_t)gridDim.x * blockDim.x)
        dst[i] = (float)src[i];
    }
  }
}

// uint8 rows copied as they are (the tensor-core training kernels read uint8 counts): out[b, :] = X[idx[b], :].
// With X in pinned host memory (mapped) this is the H2D copy AND the row gather of a host-streamed minibatch.
template <bool VEC>
__global__ void gather_rows_u8_raw_kernel(const uint8_t* __restrict__ X, int64_t ldx, const int64_t* __restrict__ idx,
                                          int64_t B, int64_t G, uint8_t* __restrict__ out, int64_t ldo) {
  for (int64_t b = blockIdx.y; b < B; b += gridDim.y) {
    const uint8_t* src = X + (idx ? idx[b] : b) * ldx;
    uint8_t* dst = out + b * ldo;
    if (VEC) {
      const int64_t nv = G >> 4;
      for (int64_t i = blockIdx.x * blockDim.x + threadIdx.x; i < nv; i += (int64_t)gridDim.x * blockDim.x) {
        uint4 raw;
        asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                     : "=r"(raw.x), "=r"(raw.y), "=r"(raw.z), "=r"(raw.w)
                     : "l"(src + 16 * i));
        *reinterpret_cast<uint4*>(dst + 16 * i) = raw;
      }
      for (int64_t i = nv * 16 + blockIdx.x * blockDim.x + threadIdx.x; i < G; i += (int64_t)gridDim.x * blockDim.x)
        dst[i] = src[i];
    } else {
      for (int64_t i = blockIdx.x * blockDim.x + threadIdx.x; i < G; i += (int64_t)gridDim.x * blockDim.x)
        dst[i] = src[i];
    }
  }
}

template <typename T>
static int launch_gather_int(const T* X, int64_t ldx, const int64_t* idx, int64_t B, int64_t G, float* out,
                             int64_t ldo, cudaStream_t st, const char* name) {
  constexpr int PER = 16 / (int)sizeof(T);
  const bool vec = (ldx % PER == 0) && ((uintptr_t)X % 16 == 0) && (ldo % 4 == 0) && ((uintptr_t)out % 16 == 0) &&
                   G >= PER;
  dim3 grid((unsigned)std::min<int64_t>(cdiv(G, (vec ? PER : 1) * 2 * 256), 64), (unsigned)std::min<int64_t>(B, 65535));
  if (vec)
    gather_rows_int_kernel<T, true><<<grid, 256, 0, st>>>(X, ldx, idx, B, G, out, ldo);
  else
    gather_rows_int_kernel<T, false><<<grid, 256, 0, st>>>(X, ldx, idx, B, G, out, ldo);
  SCVI_LAUNCH_CHECK(name);
  return SCVI_OK;
}

// CSR-resident counts: out[b, :] = dense row (idx ? idx[b] : b) of the sparse matrix -- the reference's
// `data[indexes].toarray().astype(float32)` (src/scvi/data/_anntorchdataset.py:143-152).  One CTA per row:
// zero-fill with 16-byte stores, barrier, scatter the row's non-zeros.  IT = uint16_t / int32_t column ids,
// VT = float / uint16_t / uint8_t values.  Column ids >= G are skipped and counted in *n_bad (host raises).
template <typename IT, typename VT>
__global__ void gather_rows_csr_kernel(const int64_t* __restrict__ indptr, const IT* __restrict__ indices,
                                       const VT* __restrict__ values, const int64_t* __restrict__ idx, int64_t B,
                                       int64_t G, float* __restrict__ out, int64_t ldo, int vec,
                                       unsigned long long* __restrict__ n_bad) {
  for (int64_t b = blockIdx.x; b < B; b += gridDim.x) {
    const int64_t row = idx ? idx[b] : b;
    float* dst = out + b * ldo;
    if (vec) {
      const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int64_t i = threadIdx.x; i < (G >> 2); i += blockDim.x) reinterpret_cast<float4*>(dst)[i] = z;
      for (int64_t i = (G & ~(int64_t)3) + threadIdx.x; i < G; i += blockDim.x) dst[i] = 0.f;
    } else {
      for (int64_t i = threadIdx.x; i < G; i += blockDim.x) dst[i] = 0.f;
    }
    __syncthreads();
    const int64_t p0 = indptr[row], p1 = indptr[row + 1];
    for (int64_t p = p0 + threadIdx.x; p < p1; p += blockDim.x) {
      const int64_t c = (int64_t)indices[p];
      if (c >= 0 && c < G)
        dst[c] = (float)values[p];
      else if (n_bad)
        atomicAdd(n_bad, 1ull);
    }
    __syncthreads();  // the next row of this CTA may alias nothing, but keep the phases separate
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
  // one CTA per GR_ROWS rows (2 x 4 x 16 B in flight per thread at G=2000); very long rows split over grid.x
  dim3 grid((unsigned)std::min<int64_t>(cdiv(per_row, 16 * threads), 64),
            (unsigned)std::min<int64_t>(cdiv(B, GR_ROWS), 65535));
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
  SCVI_REQUIRE(X && out, "scvi_gather_rows_u16: NULL pointer");
  SCVI_REQUIRE(n_rows >= 0 && G >= 0 && B >= 0 && ldx >= G && ldo >= G, "scvi_gather_rows_u16: bad shape");
  SCVI_REQUIRE(idx || B <= n_rows, "scvi_gather_rows_u16: identity expansion of more rows than the matrix holds");
  if (B == 0 || G == 0) return SCVI_OK;
  return launch_gather_int<uint16_t>(X, ldx, idx, B, G, out, ldo, (cudaStream_t)stream, "scvi_gather_rows_u16");
}

int scvi_gather_rows_u8_raw(const uint8_t* X, int64_t n_rows, int64_t G, int64_t ldx, const int64_t* idx, int64_t B,
                            uint8_t* out, int64_t ldo, void* stream) {
  using namespace scvi;
  SCVI_REQUIRE(X && out, "scvi_gather_rows_u8_raw: NULL pointer");
  SCVI_REQUIRE(n_rows >= 0 && G >= 0 && B >= 0 && ldx >= G && ldo >= G, "scvi_gather_rows_u8_raw: bad shape");
  SCVI_REQUIRE(idx || B <= n_rows, "scvi_gather_rows_u8_raw: identity copy of more rows than the matrix holds");
  if (B == 0 || G == 0) return SCVI_OK;
  const bool vec = (ldx % 16 == 0) && ((uintptr_t)X % 16 == 0) && (ldo % 16 == 0) && ((uintptr_t)out % 16 == 0) && G >= 16;
  dim3 grid(1, (unsigned)std::min<int64_t>(B, 65535));
  if (vec) gather_rows_u8_raw_kernel<true><<<grid, 128, 0, (cudaStream_t)stream>>>(X, ldx, idx, B, G, out, ldo);
  else gather_rows_u8_raw_kernel<false><<<grid, 128, 0, (cudaStream_t)stream>>>(X, ldx, idx, B, G, out, ldo);
  SCVI_LAUNCH_CHECK("scvi_gather_rows_u8_raw");
  return SCVI_OK;
}

int scvi_gather_rows_u8(const uint8_t* X, int64_t n_rows, int64_t G, int64_t ldx, const int64_t* idx,
                        int64_t B, float* out, int64_t ldo, void* stream) {
  using namespace scvi;
  SCVI_REQUIRE(X && out, "scvi_gather_rows_u8: NULL pointer");
  SCVI_REQUIRE(n_rows >= 0 && G >= 0 && B >= 0 && ldx >= G && ldo >= G, "scvi_gather_rows_u8: bad shape");
  SCVI_REQUIRE(idx || B <= n_rows, "scvi_gather_rows_u8: identity expansion of more rows than the matrix holds");
  if (B == 0 || G == 0) return SCVI_OK;
  return launch_gather_int<uint8_t>(X, ldx, idx, B, G, out, ldo, (cudaStream_t)stream, "scvi_gather_rows_u8");
}

int scvi_gather_rows_csr(const int64_t* indptr, const void* indices, int index_bytes, const void* values,
                         int value_kind, int64_t n_rows, int64_t G, const int64_t* idx, int64_t B, float* out,
                         int64_t ldo, unsigned long long* n_bad, void* stream) {
  using namespace scvi;
  SCVI_REQUIRE(indptr && out, "scvi_gather_rows_csr: NULL pointer");
  SCVI_REQUIRE(index_bytes == 2 || index_bytes == 4, "scvi_gather_rows_csr: index_bytes must be 2 (uint16) or 4 (int32)");
  SCVI_REQUIRE(value_kind >= 0 && value_kind <= 2, "scvi_gather_rows_csr: value_kind must be 0 (f32), 1 (u16) or 2 (u8)");
  SCVI_REQUIRE(n_rows >= 0 && G >= 0 && B >= 0 && ldo >= G, "scvi_gather_rows_csr: bad shape");
  SCVI_REQUIRE(idx || B <= n_rows, "scvi_gather_rows_csr: identity expansion of more rows than the matrix holds");
  if (B == 0 || G == 0) return SCVI_OK;
  const int vec = (ldo % 4 == 0) && ((uintptr_t)out % 16 == 0);
  const unsigned grid = (unsigned)std::min<int64_t>(B, 148 * 16);
  cudaStream_t st = (cudaStream_t)stream;
#define SCVI_CSR_LAUNCH(IT, VT)                                                                                  \
  gather_rows_csr_kernel<IT, VT><<<grid, 256, 0, st>>>(indptr, (const IT*)indices, (const VT*)values, idx, B, G, \
                                                       out, ldo, vec, n_bad)
  if (index_bytes == 2) {
    if (value_kind == 0) SCVI_CSR_LAUNCH(uint16_t, float);
    else if (value_kind == 1) SCVI_CSR_LAUNCH(uint16_t, uint16_t);
    else SCVI_CSR_LAUNCH(uint16_t, uint8_t);
  } else {
    if (value_kind == 0) SCVI_CSR_LAUNCH(int32_t, float);
    else if (value_kind == 1) SCVI_CSR_LAUNCH(int32_t, uint16_t);
    else SCVI_CSR_LAUNCH(int32_t, uint8_t);
  }
#undef SCVI_CSR_LAUNCH
  SCVI_LAUNCH_CHECK("scvi_gather_rows_csr");
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
