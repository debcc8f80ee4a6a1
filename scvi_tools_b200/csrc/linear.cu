// FCLayers linear layer: y = f(x) W[:, :K]^T + sum_c W[:, cov_col_c] + b, fp32 FFMA tiles (software-pipelined;
// split reductions for few-cell minibatches; covariate-column and bias gradients inside the dW GEMM).
// Replaces `torch.cat((x, one_hot(cat)))` + nn.Linear of FCLayers
// (reference src/scvi/nn/_base_components.py:148-188,198-215) and the encoder pre-ops
// log1p / library (src/scvi/module/_vae.py:370-374).  The one-hot concat is a weight-column
// gather in forward and a segmented column sum in backward.
#include "common.cuh"

namespace scvi {

constexpr int LT = 64;        // output tile edge (rows and cols)
constexpr int LK = 32;        // reduction chunk
constexpr int LPAD = LK + 4;  // smem row stride (floats): conflict-free float4 reads for strided rows
constexpr int LTHREADS = 256;

// ---------------------------------------------------------------------------------------
// forward
// ---------------------------------------------------------------------------------------
// epilogue of one 64 x 64 output tile held as acc[4][4] per thread (rows ty + 16 i, columns tx + 16 j):
// bias + covariate columns (the one-hot GEMM as a column gather), store, per-tile BatchNorm statistics.
// `smem`: at least LT * (LT + 1) floats, free to overwrite.
__device__ __forceinline__ void linear_fwd_epilogue(float (&acc)[4][4], float* smem, int tid, int64_t r0, int64_t n0,
                                                    int64_t B, int64_t N, const float* __restrict__ W, int64_t ldw,
                                                    const float* __restrict__ bias,
                                                    const int32_t* __restrict__ cov_cols, int n_cov,
                                                    float* __restrict__ y, int64_t ldy,
                                                    float* __restrict__ bn_partials) {
  const int tx = tid & 15, ty = tid >> 4;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int64_t r = r0 + ty + 16 * i;
    if (r >= B) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int64_t n = n0 + tx + 16 * j;
      if (n >= N) continue;
      float v = acc[i][j];
      for (int c = 0; c < n_cov; ++c) v += W[n * ldw + cov_cols[r * n_cov + c]];
      if (bias) v += bias[n];
      acc[i][j] = v;
      y[r * ldy + n] = v;
    }
  }
  // per-tile BatchNorm statistics (mean, M2) over the valid rows of this tile
  if (bn_partials != nullptr) {
    float* ts = smem;  // [LT][LT+1] staging
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) ts[(ty + 16 * i) * (LT + 1) + tx + 16 * j] = acc[i][j];
    __syncthreads();
    if (tid < LT) {
      const int64_t n = n0 + tid;
      const int cnt = (int)min((int64_t)LT, B - r0);
      if (n < N) {
        float s = 0.f;
        for (int r = 0; r < cnt; ++r) s += ts[r * (LT + 1) + tid];
        const float mean = s / (float)cnt;
        float m2 = 0.f;
        for (int r = 0; r < cnt; ++r) {
          const float d = ts[r * (LT + 1) + tid] - mean;
          m2 = fmaf(d, d, m2);
        }
        bn_partials[((int64_t)blockIdx.x * 2 + 0) * N + n] = mean;
        bn_partials[((int64_t)blockIdx.x * 2 + 1) * N + n] = m2;
      }
    }
  }
}

// SPLIT: blockIdx.z owns the reduction range [z * k_per_split, +k_per_split) and leaves its raw partial tile in
// part[z][B][N] (no epilogue); linear_fwd_finish_kernel sums the partials in z order and runs the epilogue.  For
// few-cell minibatches with a long reduction (TOTALVI's encoder: 256 cells x 4100 inputs -> 256) the 64 x 64 tiles
// alone would occupy a tenth of the SMs.
template <bool VEC, bool SPLIT>
__global__ void __launch_bounds__(LTHREADS)
    linear_fwd_kernel(const float* __restrict__ x, int64_t ldx, const int64_t* __restrict__ row_idx, int64_t B,
                      int64_t K, int64_t N, const float* __restrict__ W, int64_t ldw,
                      const float* __restrict__ bias, const int32_t* __restrict__ cov_cols, int n_cov,
                      int flags, float* __restrict__ y, int64_t ldy, float* __restrict__ library,
                      float* __restrict__ bn_partials, int64_t k_per_split, float* __restrict__ part) {
  __shared__ __align__(16) float smem[2 * LT * LPAD + LT];
  float* xs = smem;              // [LT][LPAD]
  float* ws = smem + LT * LPAD;  // [LT][LPAD]
  __shared__ int64_t rows[LT];

  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const int64_t r0 = (int64_t)blockIdx.x * LT;
  const int64_t n0 = (int64_t)blockIdx.y * LT;
  const bool log1p_in = flags & SCVI_LIN_LOG1P;
  const bool want_lib = (flags & SCVI_LIN_LIBRARY) && blockIdx.y == 0 && library != nullptr;

  if (tid < LT) {
    const int64_t r = r0 + tid;
    rows[tid] = (r < B) ? (row_idx ? row_idx[r] : r) : -1;
  }
  __syncthreads();

  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  float lib_part[2] = {0.f, 0.f};

  const int64_t k_begin = SPLIT ? (int64_t)blockIdx.z * k_per_split : 0;
  const int64_t k_end = SPLIT ? min(K, k_begin + k_per_split) : K;
  // vector path, software-pipelined: the global loads of chunk i + 1 are in flight while chunk i is multiplied out of
  // shared memory (64 rows x 8 float4 per operand; thread -> (row = tid/8 + 32*it, kq = tid%8))
  float4 xv[2], wv[2];
  const int kq = tid & 7;
  auto gload = [&](int64_t k0) {
#pragma unroll
    for (int it = 0; it < 2; ++it) {
      const int rr = (tid >> 3) + 32 * it;
      const int64_t k = k0 + 4 * kq;
      const int64_t src = rows[rr];
      const int64_t n = n0 + rr;
      xv[it] = (src >= 0 && k < K) ? ld_stream4(x + src * ldx + k) : make_float4(0.f, 0.f, 0.f, 0.f);
      wv[it] = (n < N && k < K) ? *reinterpret_cast<const float4*>(W + n * ldw + k) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
  };
  if (VEC && k_begin < k_end) gload(k_begin);
  for (int64_t k0 = k_begin; k0 < k_end; k0 += LK) {
    if (VEC) {
#pragma unroll
      for (int it = 0; it < 2; ++it) {
        const int rr = (tid >> 3) + 32 * it;
        float4 v = xv[it];
        lib_part[it] += (v.x + v.y) + (v.z + v.w);
        if (log1p_in) v = make_float4(log1pf(v.x), log1pf(v.y), log1pf(v.z), log1pf(v.w));
        *reinterpret_cast<float4*>(&xs[rr * LPAD + 4 * kq]) = v;
        *reinterpret_cast<float4*>(&ws[rr * LPAD + 4 * kq]) = wv[it];
      }
    } else {
      // scalar path: thread -> (row = tid/32 + 8*it, k = tid%32)
      const int kk = tid & 31;
#pragma unroll
      for (int it = 0; it < 8; ++it) {
        const int rr = (tid >> 5) + 8 * it;
        const int64_t k = k0 + kk;
        float v = 0.f;
        const int64_t src = rows[rr];
        if (src >= 0 && k < K) v = x[src * ldx + k];
        const float raw = v;
        if (log1p_in) v = log1pf(v);
        xs[rr * LPAD + kk] = v;
        const int64_t n = n0 + rr;
        ws[rr * LPAD + kk] = (n < N && k < K) ? W[n * ldw + k] : 0.f;
        if (want_lib) {
          // warp-level row sum of this chunk (all 32 lanes of the warp share row rr)
          float s = warp_sum(raw);
          if (kk == 0) smem[2 * LT * LPAD + rr] = (k0 == 0 ? 0.f : smem[2 * LT * LPAD + rr]) + s;
        }
      }
    }
    __syncthreads();
    if (VEC && k0 + LK < k_end) gload(k0 + LK);
    auto kstep = [&](int k) {
      float4 a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = *reinterpret_cast<const float4*>(&xs[(ty + 16 * i) * LPAD + k]);
#pragma unroll
      for (int j = 0; j < 4; ++j) b[j] = *reinterpret_cast<const float4*>(&ws[(tx + 16 * j) * LPAD + k]);
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          acc[i][j] = fmaf(a[i].x, b[j].x, acc[i][j]);
          acc[i][j] = fmaf(a[i].y, b[j].y, acc[i][j]);
          acc[i][j] = fmaf(a[i].z, b[j].z, acc[i][j]);
          acc[i][j] = fmaf(a[i].w, b[j].w, acc[i][j]);
        }
    };
    if (k_end - k0 >= LK) {
#pragma unroll
      for (int k = 0; k < LK; k += 4) kstep(k);
    } else {
      // short tail chunk (the whole reduction for the 10-wide latent inputs): the staged zeros beyond it add nothing
      const int kt = (int)(k_end - k0);
      for (int k = 0; k < kt; k += 4) kstep(k);
    }
    __syncthreads();
  }

  if (SPLIT) {
    float* pz = part + (int64_t)blockIdx.z * B * N;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int64_t r = r0 + ty + 16 * i;
      if (r >= B) continue;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int64_t n = n0 + tx + 16 * j;
        if (n < N) pz[r * N + n] = acc[i][j];
      }
    }
    return;
  }

  // library = log(sum_g x)   (reference _vae.py:372)
  if (want_lib) {
    if (VEC) {
#pragma unroll
      for (int it = 0; it < 2; ++it) {
        float s = lib_part[it];
        s += __shfl_xor_sync(0xffffffffu, s, 1);
        s += __shfl_xor_sync(0xffffffffu, s, 2);
        s += __shfl_xor_sync(0xffffffffu, s, 4);
        const int rr = (tid >> 3) + 32 * it;
        if ((tid & 7) == 0 && r0 + rr < B) library[r0 + rr] = logf(s);
      }
    } else {
      if (tid < LT && r0 + tid < B) library[r0 + tid] = logf(smem[2 * LT * LPAD + tid]);
    }
  }

  linear_fwd_epilogue(acc, smem, tid, r0, n0, B, N, W, ldw, bias, cov_cols, n_cov, y, ldy, bn_partials);
}

// sums the split-K partial tiles in split order, then the common epilogue
__global__ void __launch_bounds__(LTHREADS)
    linear_fwd_finish_kernel(const float* __restrict__ part, int n_split, int64_t B, int64_t N,
                             const float* __restrict__ W, int64_t ldw, const float* __restrict__ bias,
                             const int32_t* __restrict__ cov_cols, int n_cov, float* __restrict__ y, int64_t ldy,
                             float* __restrict__ bn_partials) {
  __shared__ __align__(16) float smem[LT * (LT + 1)];
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const int64_t r0 = (int64_t)blockIdx.x * LT;
  const int64_t n0 = (int64_t)blockIdx.y * LT;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int64_t r = r0 + ty + 16 * i, n = n0 + tx + 16 * j;
      float v = 0.f;
      if (r < B && n < N)
        for (int z = 0; z < n_split; ++z) v += part[((int64_t)z * B + r) * N + n];
      acc[i][j] = v;
    }
  linear_fwd_epilogue(acc, smem, tid, r0, n0, B, N, W, ldw, bias, cov_cols, n_cov, y, ldy, bn_partials);
}

// ---------------------------------------------------------------------------------------
// backward: dW[n, k] = sum_b dy[b, n] * f(x[b, k]), split over B, partial slabs
// ---------------------------------------------------------------------------------------
constexpr int BB = 32;  // rows per reduction chunk

// The covariate columns and the bias ride along as extra columns of the x operand: column k in [K, n_w_cols) is the
// one-hot indicator "some cov_cols[b, .] == k", column n_w_cols (when k_ext > n_w_cols) is all ones, so that
// dW[:, K:] (the one-hot part of the weight) and db come out of the same GEMM instead of three more launches.
// db_direct (non-NULL only for an unsplit launch writing straight into dW) receives the ones-column.
template <bool VEC>
__global__ void __launch_bounds__(LTHREADS)
    linear_dw_kernel(const float* __restrict__ x, int64_t ldx, const int64_t* __restrict__ row_idx, int64_t B,
                     int64_t K, int64_t N, const float* __restrict__ dy, int64_t lddy, int log1p_in,
                     int64_t rows_per_split, float* __restrict__ out, int64_t ldo, int64_t slab_stride,
                     const int32_t* __restrict__ cov_cols, int n_cov, int64_t n_w_cols, int64_t k_ext,
                     float* __restrict__ db_direct) {
  __shared__ __align__(16) float dys[BB][LT];  // [b][n]
  __shared__ __align__(16) float xs[BB][LT];   // [b][k]
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const int64_t k0 = (int64_t)blockIdx.x * LT;
  const int64_t n0 = (int64_t)blockIdx.y * LT;
  const int64_t b_begin = (int64_t)blockIdx.z * rows_per_split;
  const int64_t b_end = min(B, b_begin + rows_per_split);

  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  // software-pipelined: the loads of the next 32-row chunk are in flight while this one is multiplied
  float4 dreg[2], vreg[2];
  auto gload = [&](int64_t b0) {
    // 32 rows x 64 cols per operand = 512 float4 -> 2 per thread
#pragma unroll
    for (int it = 0; it < 2; ++it) {
      const int e = tid + LTHREADS * it;  // 0..511
      const int rr = e >> 4;              // 0..31
      const int c4 = (e & 15) * 4;
      const int64_t b = b0 + rr;
      const bool rv = b < b_end;
      float4 d = make_float4(0.f, 0.f, 0.f, 0.f), v = d;
      if (rv) {
        const int64_t src = row_idx ? row_idx[b] : b;
        if (VEC) {
          if (n0 + c4 < N) d = *reinterpret_cast<const float4*>(dy + b * lddy + n0 + c4);
          if (k0 + c4 < K) v = ld_stream4(x + src * ldx + k0 + c4);
        } else {
          float* dp = &d.x;
          float* vp = &v.x;
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            if (n0 + c4 + q < N) dp[q] = dy[b * lddy + n0 + c4 + q];
            if (k0 + c4 + q < K) vp[q] = x[src * ldx + k0 + c4 + q];
          }
        }
        if (log1p_in) v = make_float4(log1pf(v.x), log1pf(v.y), log1pf(v.z), log1pf(v.w));
        if (k0 + c4 + 3 >= K && k_ext > K) {  // this piece reaches into the one-hot / ones columns
          float* vp = &v.x;
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const int64_t k = k0 + c4 + q;
            if (k < K || k >= k_ext) continue;
            float e = 0.f;
            if (k < n_w_cols) {
              for (int c = 0; c < n_cov; ++c) e = (cov_cols[b * n_cov + c] == (int32_t)k) ? 1.f : e;
            } else {
              e = 1.f;
            }
            vp[q] = e;
          }
        }
      }
      dreg[it] = d;
      vreg[it] = v;
    }
  };
  if (b_begin < b_end) gload(b_begin);
  for (int64_t b0 = b_begin; b0 < b_end; b0 += BB) {
#pragma unroll
    for (int it = 0; it < 2; ++it) {
      const int e = tid + LTHREADS * it;
      *reinterpret_cast<float4*>(&dys[e >> 4][(e & 15) * 4]) = dreg[it];
      *reinterpret_cast<float4*>(&xs[e >> 4][(e & 15) * 4]) = vreg[it];
    }
    __syncthreads();
    if (b0 + BB < b_end) gload(b0 + BB);
#pragma unroll 8
    for (int r = 0; r < BB; ++r) {
      const float4 a = *reinterpret_cast<const float4*>(&dys[r][4 * ty]);
      const float4 b = *reinterpret_cast<const float4*>(&xs[r][4 * tx]);
      const float av[4] = {a.x, a.y, a.z, a.w};
      const float bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }
  float* o = out + (int64_t)blockIdx.z * slab_stride;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int64_t n = n0 + 4 * ty + i;
    if (n >= N) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int64_t k = k0 + 4 * tx + j;
      if (k >= k_ext) continue;
      if (db_direct != nullptr && k == n_w_cols)
        db_direct[n] = acc[i][j];
      else
        o[n * ldo + k] = acc[i][j];
    }
  }
}

// sum of the dW slabs in split order; columns [0, n_w_cols) -> dW (row stride ldw), column n_w_cols -> db
__global__ void linear_dw_finish_kernel(const float* __restrict__ slabs, int n_split, int64_t N, int64_t k_ext,
                                        int64_t n_w_cols, int64_t ldw, float* __restrict__ dW,
                                        float* __restrict__ db) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= N * k_ext) return;
  float t = 0.f;
  for (int z = 0; z < n_split; ++z) t += slabs[(int64_t)z * N * k_ext + e];
  const int64_t n = e / k_ext, k = e % k_ext;
  if (k < n_w_cols)
    dW[n * ldw + k] = t;
  else if (db)
    db[n] = t;
}

// covariate columns + bias: per row-chunk partial of
//   extra[c, n] = sum_{b : some cov_cols[b,.] == K + c} dy[b, n],   extra[n_extra, n] = sum_b dy[b, n]
// one thread per output column n (no conflicts), rows in chunks of COV_ROWS.
constexpr int COV_ROWS = 32;
__global__ void linear_dcov_kernel(const float* __restrict__ dy, int64_t lddy, int64_t B, int64_t K, int64_t N,
                                   const int32_t* __restrict__ cov_cols, int n_cov, int n_extra,
                                   float* __restrict__ part /*[chunks][n_extra+1][N]*/) {
  extern __shared__ float accs[];  // [(n_extra+1)][blockDim.x]
  const int n_local = threadIdx.x;
  const int64_t n = (int64_t)blockIdx.y * blockDim.x + n_local;
  for (int c = 0; c <= n_extra; ++c) accs[c * blockDim.x + n_local] = 0.f;
  const int64_t b0 = (int64_t)blockIdx.x * COV_ROWS;
  const int64_t b1 = min(B, b0 + COV_ROWS);
  if (n < N) {
    // rows in groups of 8, loads first (same accumulation order): one memory round trip per group
    for (int64_t bg = b0; bg < b1; bg += 8) {
      float dv[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int64_t b = bg + i;
        dv[i] = (b < b1) ? dy[b * lddy + n] : 0.f;  // a row past the chunk adds an exact zero
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int64_t b = min(bg + i, b1 - 1);
        const float d = dv[i];
        accs[n_extra * blockDim.x + n_local] += d;
        for (int c = 0; c < n_cov; ++c) {
          const int col = cov_cols[b * n_cov + c] - (int)K;
          accs[col * blockDim.x + n_local] += d;
        }
      }
    }
    for (int c = 0; c <= n_extra; ++c)
      part[((int64_t)blockIdx.x * (n_extra + 1) + c) * N + n] = accs[c * blockDim.x + n_local];
  }
}

// covariate columns / bias from the (column-summed) dcov partials: row [c][n] -> dW[n, K + c] or db[n]
__global__ void linear_dcov_scatter_kernel(const float* __restrict__ tot, int n_extra, int64_t N, int64_t K,
                                           int64_t ldw, float* __restrict__ dW, float* __restrict__ db) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= (int64_t)(n_extra + 1) * N) return;
  const int c = (int)(e / N);
  const int64_t n = e % N;
  if (c < n_extra) {
    if (dW) dW[n * ldw + K + c] = tot[e];
  } else if (db) {
    db[n] = tot[e];
  }
}
// dx[b, k] = sum_n dy[b, n] * W[n, k]   (k < K).  gridDim.z > 1: blockIdx.z reduces over its own range of
// n_per_split outputs into slab z of `dx` (slab stride B * lddx); the caller sums the slabs in order.
__global__ void __launch_bounds__(LTHREADS)
    linear_dx_kernel(const float* __restrict__ dy, int64_t lddy, int64_t B, int64_t K, int64_t N,
                     const float* __restrict__ W, int64_t ldw, float* __restrict__ dx, int64_t lddx,
                     int64_t n_per_split) {
  __shared__ float dys[LT][LK + 1];  // [b][n]
  __shared__ float wsm[LK][LT + 1];  // [n][k]
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const int64_t r0 = (int64_t)blockIdx.x * LT;
  const int64_t k0 = (int64_t)blockIdx.y * LT;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  const int64_t n_begin = (int64_t)blockIdx.z * n_per_split;
  const int64_t n_end = min(N, n_begin + n_per_split);
  dx += (int64_t)blockIdx.z * B * lddx;
  // software-pipelined: the loads of the next 32-output chunk are in flight while this one is multiplied
  float dyr[8], wr[8];
  auto gload = [&](int64_t nn0) {
#pragma unroll
    for (int it = 0; it < 8; ++it) {
      const int e = tid + LTHREADS * it;  // 0..2047
      {  // dy tile: [64 b][32 n]
        const int rr = e >> 5, nn = e & 31;
        const int64_t b = r0 + rr, n = nn0 + nn;
        dyr[it] = (b < B && n < n_end) ? dy[b * lddy + n] : 0.f;
      }
      {  // W tile: [32 n][64 k]
        const int nn = e >> 6, kk = e & 63;
        const int64_t n = nn0 + nn, k = k0 + kk;
        wr[it] = (n < n_end && k < K) ? W[n * ldw + k] : 0.f;
      }
    }
  };
  if (n_begin < n_end) gload(n_begin);
  for (int64_t nn0 = n_begin; nn0 < n_end; nn0 += LK) {
#pragma unroll
    for (int it = 0; it < 8; ++it) {
      const int e = tid + LTHREADS * it;
      dys[e >> 5][e & 31] = dyr[it];
      wsm[e >> 6][e & 63] = wr[it];
    }
    __syncthreads();
    if (nn0 + LK < n_end) gload(nn0 + LK);
#pragma unroll 8
    for (int nn = 0; nn < LK; ++nn) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = dys[ty + 16 * i][nn];
#pragma unroll
      for (int j = 0; j < 4; ++j) b[j] = wsm[nn][tx + 16 * j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int64_t b = r0 + ty + 16 * i;
    if (b >= B) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int64_t k = k0 + tx + 16 * j;
      if (k < K) dx[b * lddx + k] = acc[i][j];
    }
  }
}

// split-K plan of the forward GEMM: only when the 64 x 64 tiles alone would leave most SMs idle and the reduction
// is long enough to share out.  Returns the number of splits (1 = none) and the (LK-aligned) range per split.
static int64_t fwd_splits(int64_t B, int64_t K, int64_t N, int flags, int64_t* k_per_split) {
  const int64_t tiles = cdiv(B, LT) * cdiv(N, LT);
  const int64_t chunks = cdiv(K, LK);
  *k_per_split = K;
  if ((flags & SCVI_LIN_LIBRARY) || tiles >= 74 || chunks < 16) return 1;
  int64_t s = std::min<int64_t>(std::min<int64_t>(cdiv(2 * 148, tiles), chunks / 4), 32);
  if (s <= 1) return 1;
  const int64_t kps = cdiv(chunks, s) * LK;
  *k_per_split = kps;
  return cdiv(K, kps);
}

// the same for dx = dy W (reduction over the N outputs)
static int64_t dx_splits(int64_t B, int64_t K, int64_t N, int64_t* n_per_split) {
  const int64_t tiles = cdiv(B, LT) * cdiv(K, LT);
  const int64_t chunks = cdiv(N, LK);
  *n_per_split = N;
  if (tiles >= 74 || chunks < 16) return 1;
  int64_t s = std::min<int64_t>(std::min<int64_t>(cdiv(2 * 148, tiles), chunks / 4), 32);
  if (s <= 1) return 1;
  const int64_t nps = cdiv(chunks, s) * LK;
  *n_per_split = nps;
  return cdiv(N, nps);
}

static int64_t dw_splits(int64_t B, int64_t K, int64_t N) {
  const int64_t tiles = cdiv(K, LT) * cdiv(N, LT);
  int64_t s = cdiv(4 * 148, tiles);                  // aim for ~4 CTAs per SM
  s = std::min<int64_t>(s, cdiv(B, 4 * BB));         // at least 128 rows per split
  return std::max<int64_t>(1, std::min<int64_t>(s, 64));
}

int64_t tc_enc_fwd_ws_bytes(int64_t B, int64_t K);
int tc_enc_fwd(const void* x, int64_t ldx, const int64_t* row_idx, int64_t B, int64_t K, const float* W, int64_t ldw,
               const float* bias, int flags, float* y, int64_t ldy, float* library, float* bn_partials, void* ws,
               cudaStream_t st);
int64_t tc_enc_dw_ws_bytes(int64_t B, int64_t K);
int tc_enc_dw(const void* x, int64_t ldx, const int64_t* row_idx, int64_t B, int64_t K, const float* dy, int64_t lddy,
              int flags, float* dW, const float* dy_absmax_part, int64_t n_absmax_part, void* ws, cudaStream_t st);
}  // namespace scvi

extern "C" {

int64_t scvi_linear_fwd_ws_bytes(int64_t B, int64_t K, int64_t N, int flags) {
  int64_t kps;
  const int64_t S = scvi::fwd_splits(B, K, N, flags, &kps);
  const int64_t split = S > 1 ? 4 * S * B * N : 0;
  return std::max<int64_t>(((flags & SCVI_LIN_TC) && N == 128) ? scvi::tc_enc_fwd_ws_bytes(B, K) : 0, split);
}

int scvi_linear_fwd(const void* xv, int64_t ldx, const int64_t* row_idx, int64_t B, int64_t K, int64_t N,
                    const float* W, int64_t ldw, const float* bias, const int32_t* cov_cols, int n_cov, int flags,
                    float* y, int64_t ldy, float* library, float* bn_partials, void* ws, void* stream) {
  using namespace scvi;
  const float* x = (const float*)xv;
  SCVI_REQUIRE(x && W && y, "scvi_linear_fwd: NULL pointer");
  SCVI_REQUIRE(B >= 0 && K > 0 && N > 0 && ldx >= K && ldw >= K && ldy >= N, "scvi_linear_fwd: bad shape");
  SCVI_REQUIRE(n_cov == 0 || cov_cols, "scvi_linear_fwd: n_cov > 0 but cov_cols is NULL");
  SCVI_REQUIRE(!(flags & SCVI_LIN_LIBRARY) || library, "scvi_linear_fwd: SCVI_LIN_LIBRARY needs `library`");
  if (B == 0) return SCVI_OK;
  if (flags & SCVI_LIN_TC) {
    if (N != 128 || n_cov != 0 || !(flags & SCVI_LIN_LOG1P) || ldy % 4 != 0) {
      set_error("scvi_linear_fwd(SCVI_LIN_TC): needs N == 128, SCVI_LIN_LOG1P, no covariates, ldy %% 4 == 0");
      return SCVI_EUNSUPPORTED;
    }
    if (flags & SCVI_LIN_X_U8) {  // uint8 counts: 16-byte row pieces straight from the resident matrix
      if ((ldx % 16 != 0) || ((uintptr_t)x % 16 != 0)) {
        set_error("scvi_linear_fwd(SCVI_LIN_X_U8): the count rows must be 16-byte aligned (ldx %% 16 == 0)");
        return SCVI_EUNSUPPORTED;
      }
      return tc_enc_fwd(xv, ldx, row_idx, B, K, W, ldw, bias, flags, y, ldy, library, bn_partials, ws, (cudaStream_t)stream);
    }
    // the bulk-copy row gather needs 16-byte aligned rows; otherwise the FFMA kernel below does the work
    if ((K % 4 == 0) && (ldx % 4 == 0) && ((uintptr_t)x % 16 == 0))
      return tc_enc_fwd(x, ldx, row_idx, B, K, W, ldw, bias, flags, y, ldy, library, bn_partials, ws, (cudaStream_t)stream);
  }
  if (flags & SCVI_LIN_X_U8) {
    set_error("scvi_linear_fwd: SCVI_LIN_X_U8 needs SCVI_LIN_TC");
    return SCVI_EUNSUPPORTED;
  }
  const bool vec = (K % 4 == 0) && (ldx % 4 == 0) && (ldw % 4 == 0) && ((uintptr_t)x % 16 == 0) &&
                   ((uintptr_t)W % 16 == 0);
  dim3 grid((unsigned)cdiv(B, LT), (unsigned)cdiv(N, LT));
  cudaStream_t st = (cudaStream_t)stream;
  int64_t kps;
  const int64_t S = ws ? fwd_splits(B, K, N, flags, &kps) : 1;  // no workspace (a caller that did not ask): no split
  if (S > 1) {
    float* part = (float*)ws;
    dim3 sgrid(grid.x, grid.y, (unsigned)S);
    if (vec)
      linear_fwd_kernel<true, true><<<sgrid, LTHREADS, 0, st>>>(x, ldx, row_idx, B, K, N, W, ldw, bias, cov_cols, n_cov, flags, y, ldy, library, bn_partials, kps, part);
    else
      linear_fwd_kernel<false, true><<<sgrid, LTHREADS, 0, st>>>(x, ldx, row_idx, B, K, N, W, ldw, bias, cov_cols, n_cov, flags, y, ldy, library, bn_partials, kps, part);
    SCVI_LAUNCH_CHECK("scvi_linear_fwd(split)");
    linear_fwd_finish_kernel<<<grid, LTHREADS, 0, st>>>(part, (int)S, B, N, W, ldw, bias, cov_cols, n_cov, y, ldy, bn_partials);
    SCVI_LAUNCH_CHECK("scvi_linear_fwd(finish)");
    return SCVI_OK;
  }
  if (vec)
    linear_fwd_kernel<true, false><<<grid, LTHREADS, 0, st>>>(x, ldx, row_idx, B, K, N, W, ldw, bias, cov_cols, n_cov, flags, y, ldy, library, bn_partials, K, nullptr);
  else
    linear_fwd_kernel<false, false><<<grid, LTHREADS, 0, st>>>(x, ldx, row_idx, B, K, N, W, ldw, bias, cov_cols, n_cov, flags, y, ldy, library, bn_partials, K, nullptr);
  SCVI_LAUNCH_CHECK("scvi_linear_fwd");
  return SCVI_OK;
}

// everything but the dx slabs, which sit behind it
static int64_t linear_bwd_base_bytes(int64_t B, int64_t K, int64_t N, int64_t n_w_cols) {
  using namespace scvi;
  const int64_t KX = n_w_cols + 1;  // x columns + one-hot covariate columns + the ones column of the bias
  const int64_t S = dw_splits(B, KX, N);
  const int64_t n_extra = n_w_cols - K;
  const int64_t cov = 4 * (cdiv(B, COV_ROWS) + 1) * (n_extra + 1) * N + 256;
  const int64_t ffma = 4 * (S * N * KX + N * KX) + cov;
  const int64_t base = N == 128 ? std::max<int64_t>(ffma, tc_enc_dw_ws_bytes(B, K) + cov) : ffma;
  return (base + 255) / 256 * 256;
}

int64_t scvi_linear_bwd_ws_bytes(int64_t B, int64_t K, int64_t N, int64_t n_w_cols) {
  int64_t nps;
  const int64_t Sx = scvi::dx_splits(B, K, N, &nps);
  return linear_bwd_base_bytes(B, K, N, n_w_cols) + (Sx > 1 ? 4 * Sx * B * K : 0);
}

int scvi_linear_bwd(const void* xv, int64_t ldx, const int64_t* row_idx, int64_t B, int64_t K, int64_t N,
                    const float* W, int64_t ldw, int64_t n_w_cols, const int32_t* cov_cols, int n_cov, int flags,
                    const float* dy, int64_t lddy, float* dW, float* db, float* dx, int64_t lddx,
                    const float* dy_absmax_part, int64_t n_absmax_part, void* ws, void* stream) {
  using namespace scvi;
  const float* x = (const float*)xv;
  SCVI_REQUIRE(x && dy && ws, "scvi_linear_bwd: NULL pointer");
  SCVI_REQUIRE(B > 0 && K > 0 && N > 0 && ldx >= K && lddy >= N && n_w_cols >= K && ldw >= n_w_cols,
               "scvi_linear_bwd: bad shape");
  SCVI_REQUIRE(!(dx && (flags & SCVI_LIN_LOG1P)), "scvi_linear_bwd: dx is not available with SCVI_LIN_LOG1P");
  SCVI_REQUIRE(!dx || (W && lddx >= K), "scvi_linear_bwd: dx needs W");
  cudaStream_t st = (cudaStream_t)stream;
  const int n_extra = (int)(n_w_cols - K);
  float* slabs = (float*)ws;
  const int64_t KX = n_w_cols + 1;
  const int64_t S = dw_splits(B, KX, N);
  bool tc = (flags & SCVI_LIN_TC) != 0;
  if (tc && (N != 128 || n_extra != 0 || ldw != K || !(flags & SCVI_LIN_LOG1P))) {
    set_error("scvi_linear_bwd(SCVI_LIN_TC): needs N == 128, SCVI_LIN_LOG1P, no covariates, contiguous dW");
    return SCVI_EUNSUPPORTED;
  }
  const bool u8 = (flags & SCVI_LIN_X_U8) != 0;
  if (u8 && (!tc || dx || ldx % 16 != 0 || (uintptr_t)x % 16 != 0 || lddy % 4 != 0 || (uintptr_t)dy % 16 != 0)) {
    set_error("scvi_linear_bwd(SCVI_LIN_X_U8): needs SCVI_LIN_TC, no dx, 16-byte aligned count rows and dy");
    return SCVI_EUNSUPPORTED;
  }
  tc = tc && (u8 || ((K % 4 == 0) && (ldx % 4 == 0))) && ((uintptr_t)x % 16 == 0) && (lddy % 4 == 0) && ((uintptr_t)dy % 16 == 0);
  float* part = tc ? (float*)((uint8_t*)ws + tc_enc_dw_ws_bytes(B, K)) : slabs + S * N * KX + N * KX;
  bool cov_done = false;  // covariate columns and bias already produced by the dW GEMM
  if (dW && tc) {
    int rc = tc_enc_dw(xv, ldx, row_idx, B, K, dy, lddy, flags, dW, dy_absmax_part, n_absmax_part, ws, st);
    if (rc) return rc;
  } else if (dW) {
    const bool vec = (K % 4 == 0) && (N % 4 == 0) && (ldx % 4 == 0) && (lddy % 4 == 0) &&
                     ((uintptr_t)x % 16 == 0) && ((uintptr_t)dy % 16 == 0);
    SCVI_REQUIRE(n_cov == 0 || cov_cols, "scvi_linear_bwd: cov_cols is NULL");
    const int64_t k_ext = n_w_cols + (db ? 1 : 0);
    const int64_t rps = cdiv(cdiv(B, S), BB) * BB;
    dim3 grid((unsigned)cdiv(k_ext, LT), (unsigned)cdiv(N, LT), (unsigned)S);
    float* out = (S == 1) ? dW : slabs;
    const int64_t ldo = (S == 1) ? ldw : k_ext;
    float* dbd = (S == 1) ? db : nullptr;
    if (vec)
      linear_dw_kernel<true><<<grid, LTHREADS, 0, st>>>(x, ldx, row_idx, B, K, N, dy, lddy, flags & SCVI_LIN_LOG1P, rps, out, ldo, N * k_ext, cov_cols, n_cov, n_w_cols, k_ext, dbd);
    else
      linear_dw_kernel<false><<<grid, LTHREADS, 0, st>>>(x, ldx, row_idx, B, K, N, dy, lddy, flags & SCVI_LIN_LOG1P, rps, out, ldo, N * k_ext, cov_cols, n_cov, n_w_cols, k_ext, dbd);
    SCVI_LAUNCH_CHECK("scvi_linear_bwd(dW)");
    if (S > 1) {
      linear_dw_finish_kernel<<<(unsigned)cdiv(N * k_ext, 256), 256, 0, st>>>(slabs, (int)S, N, k_ext, n_w_cols, ldw, dW, db);
      SCVI_LAUNCH_CHECK("scvi_linear_bwd(dW finish)");
    }
    cov_done = true;
  }
  if (!cov_done && ((dW && n_extra > 0) || db)) {
    SCVI_REQUIRE(n_cov == 0 || cov_cols, "scvi_linear_bwd: cov_cols is NULL");
    const int64_t chunks = cdiv(B, COV_ROWS);
    const int bt = 128;
    const size_t sm = (size_t)(n_extra + 1) * bt * sizeof(float);
    SCVI_REQUIRE(sm <= 48 * 1024, "scvi_linear_bwd: too many covariate columns (%d)", n_extra);
    dim3 grid((unsigned)chunks, (unsigned)cdiv(N, bt));
    linear_dcov_kernel<<<grid, bt, sm, st>>>(dy, lddy, B, K, N, cov_cols, n_cov, n_extra, part);
    SCVI_LAUNCH_CHECK("scvi_linear_bwd(dcov)");
    const int64_t ne = (int64_t)(n_extra + 1) * N;
    float* tot = part + chunks * ne;
    int rc = launch_colsum(part, chunks, ne, ne, colsum_seg1(tot, ne), st, "scvi_linear_bwd(dcov reduce)");
    if (rc) return rc;
    linear_dcov_scatter_kernel<<<(unsigned)cdiv(ne, 256), 256, 0, st>>>(tot, n_extra, N, K, ldw, dW, db);
    SCVI_LAUNCH_CHECK("scvi_linear_bwd(dcov scatter)");
  }
  if (dx) {
    dim3 grid((unsigned)cdiv(B, LT), (unsigned)cdiv(K, LT));
    int64_t nps;
    const int64_t Sx = (lddx == K) ? dx_splits(B, K, N, &nps) : 1;  // the slab reduction writes a contiguous dx
    if (Sx > 1) {
      float* xslabs = (float*)((uint8_t*)ws + linear_bwd_base_bytes(B, K, N, n_w_cols));
      grid.z = (unsigned)Sx;
      linear_dx_kernel<<<grid, LTHREADS, 0, st>>>(dy, lddy, B, K, N, W, ldw, xslabs, K, nps);
      SCVI_LAUNCH_CHECK("scvi_linear_bwd(dx split)");
      int rc = launch_colsum(xslabs, Sx, B * K, B * K, colsum_seg1(dx, B * K), st, "scvi_linear_bwd(dx reduce)");
      if (rc) return rc;
    } else {
      linear_dx_kernel<<<grid, LTHREADS, 0, st>>>(dy, lddy, B, K, N, W, ldw, dx, lddx, N);
      SCVI_LAUNCH_CHECK("scvi_linear_bwd(dx)");
    }
  }
  return SCVI_OK;
}
}
