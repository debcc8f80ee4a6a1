// The [B, n_hidden] "middle" of the SCVI VAE step between the two O(B x G) ends, fused into a few kernels
// (n_hidden = 128): BatchNorm + ReLU + Dropout of the encoder layer, latent heads + reparameterisation + KL,
// decoder hidden layer with the batch covariate as a weight-column gather, its BatchNorm + ReLU; and the
// backward of all of it.  Replaces, for the fused training plan, the sequence
//   bn_finalize, bn_act_drop_fwd, latent_fwd, linear_fwd, bn_finalize, bn_act_drop_fwd            (forward)
//   bn_act_drop_bwd, linear_bwd, latent_bwd, bn_act_drop_bwd + their split reductions             (backward)
// i.e. reference FCLayers._build_layer (src/scvi/nn/_base_components.py:102-115), Encoder.forward heads
// (:311-317), the KL of VAE.loss (src/scvi/module/_vae.py:563-565) and FCLayers.forward of the decoder
// (:148-188).  Every op on this stretch touches ~4 MB: as separate launches they are latency bound.
//
// One CTA = one tile of MD_ROWS rows, 512 threads.  Batch statistics cross CTAs through partials (mean, M2 /
// sums): the producing kernel runs as clusters of MD_CL CTAs and writes one partial per cluster (merged through
// distributed shared memory), every CTA of the NEXT kernel merges the cluster partials redundantly in a
// fixed order (deterministic, no grid barrier, no atomics).
#include <cooperative_groups.h>

#include "common.cuh"

namespace scvi {
namespace cg = cooperative_groups;

constexpr int MD_H = 128;
constexpr int MD_ROWS = 32;            // rows per CTA (two CTAs per SM hide each other's latency)
constexpr int MD_RPT = MD_ROWS / 4;    // rows per thread: thread = (column, one of 4 row groups)
constexpr int MD_THREADS = 512;
constexpr int MD_LMAX = 16;            // n_latent <= 16
constexpr int MD_WD_LD = 81;           // decoder layer-0 weight row stride in smem (n_latent + n_batch <= 80)
// The kernels that produce per-tile column statistics run as thread-block clusters of MD_CL CTAs: rank 0 sums
// the cluster's tile partials through distributed shared memory (rank order: deterministic) and writes ONE
// partial per cluster, so the consumers merge T / 8 partials (8 MB of L2 traffic per kernel instead of 64 MB
// at B = 8192).  The grid is padded to a multiple of MD_CL; padded CTAs only take part in the cluster barriers.
constexpr int MD_CL = 8;
__host__ __device__ inline int64_t md_clusters(int64_t B) { return (B + MD_ROWS * MD_CL - 1) / (MD_ROWS * MD_CL); }

// s_cl[2][128] of every CTA of the cluster -> part[cluster][2][128] (sums).  All threads of all CTAs call it.
__device__ __forceinline__ void md_cluster_sum2(float* s_cl, float* __restrict__ part) {
  cg::cluster_group cl = cg::this_cluster();
  cl.sync();
  const int col = threadIdx.x & 127, q = threadIdx.x >> 7;
  if (cl.block_rank() == 0 && q < 2) {
    float acc = 0.f;
#pragma unroll
    for (int k = 0; k < MD_CL; ++k) acc += cl.map_shared_rank(s_cl, k)[q * MD_H + col];
    part[((int64_t)(blockIdx.x / MD_CL) * 2 + q) * MD_H + col] = acc;
  }
  cl.sync();  // the other CTAs' shared memory stays alive until rank 0 has read it
}

// s_cl = (mean[128], M2[128]) of every CTA's tile -> part[cluster][2][128] = (mean, M2) of the cluster's rows
// (count-weighted merge, rank order).  Tiles past the end of the batch have count 0.
__device__ __forceinline__ void md_cluster_stats(float* s_cl, int64_t B, float* __restrict__ part) {
  cg::cluster_group cl = cg::this_cluster();
  cl.sync();
  if (cl.block_rank() == 0 && threadIdx.x < MD_H) {
    const int col = threadIdx.x;
    const int64_t cid = blockIdx.x / MD_CL, r0c = cid * MD_CL * MD_ROWS;
    float n[MD_CL], m[MD_CL], ntot = 0.f, msum = 0.f;
#pragma unroll
    for (int k = 0; k < MD_CL; ++k) {
      n[k] = (float)max((int64_t)0, min((int64_t)MD_ROWS, B - (r0c + (int64_t)k * MD_ROWS)));
      m[k] = cl.map_shared_rank(s_cl, k)[col];
      ntot += n[k];
      msum = fmaf(n[k], m[k], msum);
    }
    const float mean = msum / ntot;
    float m2 = 0.f;
#pragma unroll
    for (int k = 0; k < MD_CL; ++k) {
      const float d = m[k] - mean;
      m2 += fmaf(n[k], d * d, cl.map_shared_rank(s_cl, k)[MD_H + col]);
    }
    part[(cid * 2 + 0) * MD_H + col] = mean;
    part[(cid * 2 + 1) * MD_H + col] = m2;
  }
  cl.sync();
}

// merge per-tile (mean, M2) partials [T][2][128] -> batch mean / biased variance of column `col`:
//   mean = sum_t n_t mean_t / B,   M2 = sum_t (M2_t + n_t (mean_t - mean)^2)
// two passes of independent loads (no serial dependence through divisions), 4 thread groups x T/4 tiles each,
// partial sums merged in a fixed order.  All 512 threads call it.
__device__ __forceinline__ void md_merge_stats(const float* __restrict__ part, int64_t B, int tile_rows,
                                               float* s_scratch /*[4][128]*/, float& mean_out, float& var_out,
                                               float& cnt_out) {
  const int col = threadIdx.x & 127, grp = threadIdx.x >> 7;
  const int64_t T = (B + tile_rows - 1) / tile_rows;
  const float tr = (float)tile_rows;
  const float c_last = (float)(B - (T - 1) * tile_rows);
  float s = 0.f;
#pragma unroll 16
  for (int64_t t = grp; t < T; t += 4) s = fmaf(t == T - 1 ? c_last : tr, part[(t * 2 + 0) * MD_H + col], s);
  s_scratch[grp * MD_H + col] = s;
  __syncthreads();
  const float mean = ((s_scratch[col] + s_scratch[MD_H + col]) + (s_scratch[2 * MD_H + col] + s_scratch[3 * MD_H + col])) / (float)B;
  __syncthreads();
  float m2 = 0.f;
#pragma unroll 16
  for (int64_t t = grp; t < T; t += 4) {
    const float d = part[(t * 2 + 0) * MD_H + col] - mean;
    m2 += fmaf(t == T - 1 ? c_last : tr, d * d, part[(t * 2 + 1) * MD_H + col]);
  }
  s_scratch[grp * MD_H + col] = m2;
  __syncthreads();
  m2 = (s_scratch[col] + s_scratch[MD_H + col]) + (s_scratch[2 * MD_H + col] + s_scratch[3 * MD_H + col]);
  __syncthreads();
  mean_out = mean;
  var_out = m2 / (float)B;
  cnt_out = (float)B;
}

// (S1, S2)[col] = sum_t part[t][0 / 1][col] (fixed order), all 512 threads
__device__ __forceinline__ void md_merge_sum2(const float* __restrict__ part, int64_t T, float* s_scratch /*[8][128]*/,
                                              float& S1, float& S2) {
  const int col = threadIdx.x & 127, grp = threadIdx.x >> 7;
  float a = 0.f, b = 0.f;
#pragma unroll 8
  for (int64_t t = grp; t < T; t += 4) {
    a += part[(t * 2 + 0) * MD_H + col];
    b += part[(t * 2 + 1) * MD_H + col];
  }
  s_scratch[grp * MD_H + col] = a;
  s_scratch[(4 + grp) * MD_H + col] = b;
  __syncthreads();
  S1 = (s_scratch[col] + s_scratch[MD_H + col]) + (s_scratch[2 * MD_H + col] + s_scratch[3 * MD_H + col]);
  S2 = (s_scratch[4 * MD_H + col] + s_scratch[5 * MD_H + col]) + (s_scratch[6 * MD_H + col] + s_scratch[7 * MD_H + col]);
  __syncthreads();
}

// per-tile (mean, M2) of a column from the 4 row groups' MD_RPT values each
__device__ __forceinline__ void md_tile_stats(const float (&vals)[MD_RPT], int rg, int col, int cnt, float* s_scratch,
                                              float* __restrict__ part_out /*[2][128] of this tile*/) {
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < MD_RPT; ++i) s += (rg * MD_RPT + i < cnt) ? vals[i] : 0.f;
  s_scratch[rg * MD_H + col] = s;
  __syncthreads();
  const float mean = ((s_scratch[col] + s_scratch[MD_H + col]) + (s_scratch[2 * MD_H + col] + s_scratch[3 * MD_H + col])) / (float)cnt;
  __syncthreads();
  float m2 = 0.f;
#pragma unroll
  for (int i = 0; i < MD_RPT; ++i)
    if (rg * MD_RPT + i < cnt) {
      const float d = vals[i] - mean;
      m2 = fmaf(d, d, m2);
    }
  s_scratch[rg * MD_H + col] = m2;
  __syncthreads();
  if (rg == 0) {
    part_out[col] = mean;
    part_out[MD_H + col] = (s_scratch[col] + s_scratch[MD_H + col]) + (s_scratch[2 * MD_H + col] + s_scratch[3 * MD_H + col]);
  }
  __syncthreads();
}

struct MidFwdArgs {
  const float *y1, *part1, *gamma1, *beta1;
  float *rm1, *rv1;
  const uint8_t* mask;
  float drop_scale;
  const float *noise, *Wm, *bm, *Wv, *bv;
  float var_eps;
  const float *Wd, *bd;
  const int32_t* cov_cols;  // [B] absolute column of Wd selected by the batch covariate, or NULL
  int ncols;                // columns of Wd (n_latent + n_batch)
  float bn_eps, momentum;
  int64_t B;
  int L;
  float *a1, *q_m, *q_v, *z, *kl, *y2, *part2, *mean1, *rstd1;
};

// forward, stage 1: BN1 + ReLU + Dropout -> latent heads, z, KL -> decoder hidden pre-activation y2 + its tile stats
__device__ __forceinline__ void mid_fwd_body(const MidFwdArgs& A, float* s_cl) {
  extern __shared__ __align__(16) float sm[];
  float* s_a = sm;                          // [64][128]
  float* s_W = s_a + MD_ROWS * MD_H;        // [2L][129]  latent heads (Wm | Wv)
  float* s_Wd = s_W + 2 * MD_LMAX * 129;    // [128][MD_WD_LD]
  float* s_q = s_Wd + MD_H * MD_WD_LD;      // [64][2L]
  float* s_z = s_q + MD_ROWS * 2 * MD_LMAX; // [64][L]
  float* s_klt = s_z + MD_ROWS * MD_LMAX;   // [64][L]
  float* s_scr = s_klt + MD_ROWS * MD_LMAX; // [12][128]
  const int tid = threadIdx.x, col = tid & 127, rg = tid >> 7;
  const int L = A.L, L2 = 2 * A.L;
  const int64_t r0 = (int64_t)blockIdx.x * MD_ROWS;
  const int cnt = (int)min((int64_t)MD_ROWS, A.B - r0);

  for (int e = tid; e < L2 * MD_H; e += MD_THREADS) {
    const int j = e / MD_H, c = e % MD_H;
    s_W[j * 129 + c] = (j < L) ? A.Wm[j * MD_H + c] : A.Wv[(j - L) * MD_H + c];
  }
  for (int e = tid; e < MD_H * A.ncols; e += MD_THREADS) s_Wd[(e / A.ncols) * MD_WD_LD + e % A.ncols] = A.Wd[e];

  float mean, var, n;
  md_merge_stats(A.part1, A.B, SCVI_BN_TILE, s_scr, mean, var, n);
  const float rstd = 1.f / sqrtf(var + A.bn_eps);
  if (blockIdx.x == 0 && rg == 0) {
    A.mean1[col] = mean;
    A.rstd1[col] = rstd;
    if (A.rm1) A.rm1[col] = (1.f - A.momentum) * A.rm1[col] + A.momentum * mean;
    if (A.rv1) A.rv1[col] = (1.f - A.momentum) * A.rv1[col] + A.momentum * (var * n / fmaxf(n - 1.f, 1.f));
  }
  const float g1 = A.gamma1[col], b1 = A.beta1[col];
  {  // all of the thread's loads first (independent, in flight together), then the math
    float yv[MD_RPT];
    uint8_t mk[MD_RPT];
#pragma unroll
    for (int i = 0; i < MD_RPT; ++i) {
      const int r = rg * MD_RPT + i;
      const int64_t o = (r0 + r) * MD_H + col;
      yv[i] = (r < cnt) ? A.y1[o] : 0.f;
      mk[i] = (r < cnt && A.mask) ? A.mask[o] : (uint8_t)1;
    }
#pragma unroll
    for (int i = 0; i < MD_RPT; ++i) {
      const int r = rg * MD_RPT + i;
      float a = 0.f;
      if (r < cnt) {
        const int64_t o = (r0 + r) * MD_H + col;
        const float u = (yv[i] - mean) * (rstd * g1) + b1;
        a = fmaxf(u, 0.f);
        if (A.mask) a = mk[i] ? a * A.drop_scale : 0.f;
        A.a1[o] = a;
      }
      s_a[r * MD_H + col] = a;
    }
  }
  __syncthreads();
  // latent heads: 64 x 2L dot products of length 128
  for (int o = tid; o < MD_ROWS * L2; o += MD_THREADS) {
    const int r = o / L2, j = o % L2;
    const float* ar = s_a + r * MD_H;
    const float* wr = s_W + j * 129;
    float s = (j < L) ? A.bm[j] : A.bv[j - L];
#pragma unroll 8
    for (int c = 0; c < MD_H; ++c) s = fmaf(ar[c], wr[c], s);
    s_q[r * L2 + j] = s;
  }
  __syncthreads();
  for (int o = tid; o < MD_ROWS * L; o += MD_THREADS) {
    const int r = o / L, l = o % L;
    float zv = 0.f, klt = 0.f;
    if (r < cnt) {
      const float m = s_q[r * L2 + l];
      const float v = expf(s_q[r * L2 + L + l]) + A.var_eps;
      const float sd = sqrtf(v);
      const float s2 = sd * sd;
      const int64_t i = (r0 + r) * L + l;
      zv = fmaf(sd, A.noise[i], m);
      A.q_m[i] = m;
      A.q_v[i] = v;
      A.z[i] = zv;
      klt = 0.5f * (s2 + m * m - 1.f - logf(s2));
    }
    s_z[r * L + l] = zv;
    s_klt[r * L + l] = klt;
  }
  __syncthreads();
  if (tid < cnt) {
    float s = 0.f;
    for (int l = 0; l < L; ++l) s += s_klt[tid * L + l];
    A.kl[r0 + tid] = s;
  }
  // decoder hidden layer: y2 = z Wd[:, :L]^T + Wd[:, cov] + bd
  const float bdv = A.bd ? A.bd[col] : 0.f;
  float vals[MD_RPT];
#pragma unroll 4
  for (int i = 0; i < MD_RPT; ++i) {
    const int r = rg * MD_RPT + i;
    float v = 0.f;
    if (r < cnt) {
      v = bdv;
      for (int l = 0; l < L; ++l) v = fmaf(s_z[r * L + l], s_Wd[col * MD_WD_LD + l], v);
      if (A.cov_cols) v += s_Wd[col * MD_WD_LD + A.cov_cols[r0 + r]];
      A.y2[(r0 + r) * MD_H + col] = v;
    }
    vals[i] = v;
  }
  md_tile_stats(vals, rg, col, cnt, s_scr, s_cl);
}

__global__ void __cluster_dims__(MD_CL, 1, 1) __launch_bounds__(MD_THREADS) mid_fwd_kernel(MidFwdArgs A) {
  __shared__ float s_cl[2 * MD_H];
  if ((int64_t)blockIdx.x * MD_ROWS < A.B) mid_fwd_body(A, s_cl);
  else if (threadIdx.x < 2 * MD_H) s_cl[threadIdx.x] = 0.f;  // a CTA that pads the grid: an empty tile
  md_cluster_stats(s_cl, A.B, A.part2);
}

// forward, stage 2: BN2 + ReLU -> h
__global__ void __launch_bounds__(MD_THREADS)
    mid_fwd2_kernel(const float* __restrict__ y2, const float* __restrict__ part2, const float* __restrict__ gamma2,
                    const float* __restrict__ beta2, float* __restrict__ rm2, float* __restrict__ rv2, float bn_eps,
                    float momentum, int64_t B, float* __restrict__ h, float* __restrict__ mean2,
                    float* __restrict__ rstd2) {
  __shared__ float s_scr[12 * MD_H];
  const int col = threadIdx.x & 127, rg = threadIdx.x >> 7;
  const int64_t r0 = (int64_t)blockIdx.x * MD_ROWS;
  const int cnt = (int)min((int64_t)MD_ROWS, B - r0);
  float mean, var, n;
  md_merge_stats(part2, B, MD_ROWS * MD_CL, s_scr, mean, var, n);
  const float rstd = 1.f / sqrtf(var + bn_eps);
  if (blockIdx.x == 0 && rg == 0) {
    mean2[col] = mean;
    rstd2[col] = rstd;
    if (rm2) rm2[col] = (1.f - momentum) * rm2[col] + momentum * mean;
    if (rv2) rv2[col] = (1.f - momentum) * rv2[col] + momentum * (var * n / fmaxf(n - 1.f, 1.f));
  }
  const float g = gamma2[col], b = beta2[col];
#pragma unroll 4
  for (int i = 0; i < MD_RPT; ++i) {
    const int r = rg * MD_RPT + i;
    if (r < cnt) {
      const int64_t o = (r0 + r) * MD_H + col;
      h[o] = fmaxf((y2[o] - mean) * (rstd * g) + b, 0.f);
    }
  }
}

// ---------------------------------------------------------------------------------------
// backward
// ---------------------------------------------------------------------------------------
// stage 0: per-tile column sums of du2 = dh * 1[u2 > 0] and du2 * xhat2  (BatchNorm-2 backward, phase 1)
__global__ void __cluster_dims__(MD_CL, 1, 1) __launch_bounds__(MD_THREADS)
    mid_bwd0_kernel(const float* __restrict__ dh, const float* __restrict__ y2, const float* __restrict__ mean2,
                    const float* __restrict__ rstd2, const float* __restrict__ gamma2, const float* __restrict__ beta2,
                    int64_t B, float* __restrict__ bpart2 /*[clusters][2][128]*/) {
  __shared__ float s_scr[8 * MD_H];
  __shared__ float s_cl[2 * MD_H];
  const int col = threadIdx.x & 127, rg = threadIdx.x >> 7;
  const int64_t r0 = (int64_t)blockIdx.x * MD_ROWS;
  const int cnt = (int)max((int64_t)0, min((int64_t)MD_ROWS, B - r0));  // 0 in the CTAs that pad the grid
  const float mu = mean2[col], rs = rstd2[col], g = gamma2[col], b = beta2[col];
  float s1 = 0.f, s2 = 0.f;
  float yv[MD_RPT], dv[MD_RPT];  // dh is loaded whatever the ReLU mask says: no load waits on another
#pragma unroll
  for (int i = 0; i < MD_RPT; ++i) {
    const int r = rg * MD_RPT + i;
    const int64_t o = (r0 + r) * MD_H + col;
    yv[i] = (r < cnt) ? y2[o] : 0.f;
    dv[i] = (r < cnt) ? dh[o] : 0.f;
  }
#pragma unroll
  for (int i = 0; i < MD_RPT; ++i) {
    const int r = rg * MD_RPT + i;
    if (r < cnt) {
      const float xhat = (yv[i] - mu) * rs;
      const float d = (xhat * g + b > 0.f) ? dv[i] : 0.f;
      s1 += d;
      s2 = fmaf(d, xhat, s2);
    }
  }
  s_scr[rg * MD_H + col] = s1;
  s_scr[(4 + rg) * MD_H + col] = s2;
  __syncthreads();
  if (rg == 0) {
    s_cl[col] = (s_scr[col] + s_scr[MD_H + col]) + (s_scr[2 * MD_H + col] + s_scr[3 * MD_H + col]);
    s_cl[MD_H + col] = (s_scr[4 * MD_H + col] + s_scr[5 * MD_H + col]) + (s_scr[6 * MD_H + col] + s_scr[7 * MD_H + col]);
  }
  md_cluster_sum2(s_cl, bpart2);
}

struct MidBwdArgs {
  const float *dh, *y2, *mean2, *rstd2, *gamma2, *beta2, *bpart2;
  const float *z, *Wd;
  const int32_t* cov_cols;
  int ncols;
  const float *g_kl, *q_m, *q_v, *noise;
  float var_eps;
  const float *Wm, *Wv, *a1, *y1, *mean1, *rstd1, *gamma1, *beta1;
  const uint8_t* mask;
  float drop_scale;
  int64_t B;
  int L;
  float *da1, *bpart1;    // [B,128], [T][2][128]
  float* ppart;           // per-tile parameter-gradient partials, record layout below
  int64_t prec;           // floats per tile record
  float *dgamma2, *dbeta2;
};
// per-tile record: [dWd 128 x ncols][dbd 128][dWm L x 128][dWv L x 128][dbm L][dbv L][dbe 128]
__host__ __device__ inline int64_t md_rec_floats(int ncols, int L) { return (int64_t)MD_H * ncols + MD_H + 2 * L * MD_H + 2 * L + MD_H; }

// stage 1: BN2 backward -> dy2 -> dWd / dbd partials, dz -> latent backward -> da1, dWm / dWv / dbm / dbv partials
//          -> BN1 backward phase 1 (per-tile sums of du1, du1 * xhat1)
__device__ __forceinline__ void mid_bwd1_body(const MidBwdArgs& A, float* s_cl) {
  extern __shared__ __align__(16) float sm[];
  float* s_t = sm;                           // [64][128]  dy2, later a1
  float* s_W = s_t + MD_ROWS * MD_H;         // [2L][129]
  float* s_Wd = s_W + 2 * MD_LMAX * 129;     // [128][MD_WD_LD]
  float* s_z = s_Wd + MD_H * MD_WD_LD;       // [64][L]
  float* s_dq = s_z + MD_ROWS * MD_LMAX;     // [64][2L]
  float* s_scr = s_dq + MD_ROWS * 2 * MD_LMAX;  // [12][128]
  __shared__ int s_cat[MD_ROWS];
  const int tid = threadIdx.x, col = tid & 127, rg = tid >> 7;
  const int L = A.L, L2 = 2 * A.L, ncols = A.ncols;
  const int64_t r0 = (int64_t)blockIdx.x * MD_ROWS;
  const int cnt = (int)min((int64_t)MD_ROWS, A.B - r0);
  float* rec = A.ppart + (int64_t)blockIdx.x * A.prec;

  for (int e = tid; e < L2 * MD_H; e += MD_THREADS) {
    const int j = e / MD_H, c = e % MD_H;
    s_W[j * 129 + c] = (j < L) ? A.Wm[j * MD_H + c] : A.Wv[(j - L) * MD_H + c];
  }
  for (int e = tid; e < MD_H * ncols; e += MD_THREADS) s_Wd[(e / ncols) * MD_WD_LD + e % ncols] = A.Wd[e];
  for (int e = tid; e < MD_ROWS * L; e += MD_THREADS) s_z[e] = (e / L < cnt) ? A.z[r0 * L + e] : 0.f;
  if (tid < MD_ROWS) s_cat[tid] = (A.cov_cols && tid < cnt) ? A.cov_cols[r0 + tid] : -1;

  // the thread's y2 / dh values: loaded ahead of the partial merge so their latency hides behind it
  float yv2[MD_RPT], dv2[MD_RPT];
#pragma unroll
  for (int i = 0; i < MD_RPT; ++i) {
    const int r = rg * MD_RPT + i;
    const int64_t o = (r0 + r) * MD_H + col;
    yv2[i] = (r < cnt) ? A.y2[o] : 0.f;
    dv2[i] = (r < cnt) ? A.dh[o] : 0.f;
  }
  float S1, S2;
  md_merge_sum2(A.bpart2, md_clusters(A.B), s_scr, S1, S2);
  if (blockIdx.x == 0 && rg == 0) {
    if (A.dbeta2) A.dbeta2[col] = S1;
    if (A.dgamma2) A.dgamma2[col] = S2;
  }
  const float invB = 1.f / (float)A.B;
  {
    const float mu = A.mean2[col], rs = A.rstd2[col], g = A.gamma2[col], b = A.beta2[col];
#pragma unroll
    for (int i = 0; i < MD_RPT; ++i) {
      const int r = rg * MD_RPT + i;
      float v = 0.f;
      if (r < cnt) {
        const float xhat = (yv2[i] - mu) * rs;
        const float d = (xhat * g + b > 0.f) ? dv2[i] : 0.f;
        v = rs * g * (d - S1 * invB - xhat * (S2 * invB));
      }
      s_t[r * MD_H + col] = v;  // dy2
    }
  }
  __syncthreads();
  // dWd[n][j] partial: thread (n = col, rg) handles columns j = rg, rg + 4, ...; j == ncols is the bias
  for (int j = rg; j <= ncols; j += 4) {
    float s = 0.f;
    if (j < L) {
      for (int r = 0; r < cnt; ++r) s = fmaf(s_t[r * MD_H + col], s_z[r * L + j], s);
    } else if (j < ncols) {
      for (int r = 0; r < cnt; ++r) s += (s_cat[r] == j) ? s_t[r * MD_H + col] : 0.f;
    } else {
      for (int r = 0; r < cnt; ++r) s += s_t[r * MD_H + col];
    }
    if (j < ncols) rec[col * ncols + j] = s;
    else rec[MD_H * ncols + col] = s;
  }
  // dz[r][l] = sum_n dy2[r][n] Wd[n][l]; then the latent backward per (r, l)
  for (int o = tid; o < MD_ROWS * L; o += MD_THREADS) {
    const int r = o / L, l = o % L;
    float dm = 0.f, dt = 0.f;
    if (r < cnt) {
      float dz = 0.f;
#pragma unroll 8
      for (int n = 0; n < MD_H; ++n) dz = fmaf(s_t[r * MD_H + n], s_Wd[n * MD_WD_LD + l], dz);
      const int64_t i = (r0 + r) * L + l;
      const float m = A.q_m[i], v = A.q_v[i];
      const float gk = A.g_kl ? A.g_kl[r0 + r] : 0.f;
      dm = dz + gk * m;
      const float sd = sqrtf(v);
      const float dv = dz * A.noise[i] / (2.f * sd) + gk * 0.5f * (1.f - 1.f / v);
      dt = dv * (v - A.var_eps);
    }
    s_dq[r * L2 + l] = dm;
    s_dq[r * L2 + L + l] = dt;
  }
  __syncthreads();
  // da1[r][c] = sum_j dq[r][j] W[j][c]; BN1 backward phase 1 on the fly
  {
    const float mu = A.mean1[col], rs = A.rstd1[col], g = A.gamma1[col], b = A.beta1[col];
    float s1 = 0.f, s2 = 0.f;
    float y1v[MD_RPT];
    uint8_t mk[MD_RPT];
#pragma unroll
    for (int i = 0; i < MD_RPT; ++i) {
      const int r = rg * MD_RPT + i;
      const int64_t o = (r0 + r) * MD_H + col;
      y1v[i] = (r < cnt) ? A.y1[o] : 0.f;
      mk[i] = (r < cnt && A.mask) ? A.mask[o] : (uint8_t)1;
    }
#pragma unroll
    for (int i = 0; i < MD_RPT; ++i) {
      const int r = rg * MD_RPT + i;
      if (r < cnt) {
        float s = 0.f;
        for (int j = 0; j < L2; ++j) s = fmaf(s_dq[r * L2 + j], s_W[j * 129 + col], s);
        const int64_t o = (r0 + r) * MD_H + col;
        A.da1[o] = s;
        const float xhat = (y1v[i] - mu) * rs;
        float d = s;
        if (A.mask) d = mk[i] ? d * A.drop_scale : 0.f;
        if (!(xhat * g + b > 0.f)) d = 0.f;
        s1 += d;
        s2 = fmaf(d, xhat, s2);
      }
    }
    __syncthreads();  // all threads are done reading dy2 from s_t
    s_scr[rg * MD_H + col] = s1;
    s_scr[(4 + rg) * MD_H + col] = s2;
  }
  // a1 tile -> s_t for the latent weight gradients
#pragma unroll
  for (int i = 0; i < MD_RPT; ++i) {
    const int r = rg * MD_RPT + i;
    s_t[r * MD_H + col] = (r < cnt) ? A.a1[(r0 + r) * MD_H + col] : 0.f;
  }
  __syncthreads();
  if (rg == 0) {
    s_cl[col] = (s_scr[col] + s_scr[MD_H + col]) + (s_scr[2 * MD_H + col] + s_scr[3 * MD_H + col]);
    s_cl[MD_H + col] = (s_scr[4 * MD_H + col] + s_scr[5 * MD_H + col]) + (s_scr[6 * MD_H + col] + s_scr[7 * MD_H + col]);
  }
  // dW latent [j][c] partial = sum_r dq[r][j] a1[r][c]
  float* recL = rec + (int64_t)MD_H * ncols + MD_H;
  for (int j = rg; j < L2; j += 4) {
    float s = 0.f;
    for (int r = 0; r < cnt; ++r) s = fmaf(s_dq[r * L2 + j], s_t[r * MD_H + col], s);
    recL[j * MD_H + col] = s;  // rows 0..L-1 = dWm, L..2L-1 = dWv (contiguous)
  }
  if (tid < L2) {
    float s = 0.f;
    for (int r = 0; r < cnt; ++r) s += s_dq[r * L2 + tid];
    recL[(int64_t)L2 * MD_H + tid] = s;  // dbm | dbv
  }
}

__global__ void __cluster_dims__(MD_CL, 1, 1) __launch_bounds__(MD_THREADS) mid_bwd1_kernel(MidBwdArgs A) {
  __shared__ float s_cl[2 * MD_H];
  if ((int64_t)blockIdx.x * MD_ROWS < A.B) mid_bwd1_body(A, s_cl);
  else if (threadIdx.x < 2 * MD_H) s_cl[threadIdx.x] = 0.f;  // a CTA that pads the grid
  md_cluster_sum2(s_cl, A.bpart1);
}

// stage 2: BN1 backward phase 2 -> dy1 (+ per-tile column sums = gradient of the encoder Linear bias)
__global__ void __launch_bounds__(MD_THREADS)
    mid_bwd2_kernel(const float* __restrict__ da1, const float* __restrict__ y1, const float* __restrict__ mean1,
                    const float* __restrict__ rstd1, const float* __restrict__ gamma1, const float* __restrict__ beta1,
                    const uint8_t* __restrict__ mask, float drop_scale, const float* __restrict__ bpart1, int64_t B,
                    float* __restrict__ dy1, float* __restrict__ dgamma1, float* __restrict__ dbeta1,
                    float* __restrict__ ppart, int64_t prec, int64_t dbe_off, float* __restrict__ absmax_part) {
  __shared__ float s_scr[8 * MD_H];
  const int col = threadIdx.x & 127, rg = threadIdx.x >> 7;
  const int64_t r0 = (int64_t)blockIdx.x * MD_ROWS;
  const int cnt = (int)min((int64_t)MD_ROWS, B - r0);
  float S1, S2;
  md_merge_sum2(bpart1, md_clusters(B), s_scr, S1, S2);
  if (blockIdx.x == 0 && rg == 0) {
    if (dbeta1) dbeta1[col] = S1;
    if (dgamma1) dgamma1[col] = S2;
  }
  const float invB = 1.f / (float)B;
  const float mu = mean1[col], rs = rstd1[col], g = gamma1[col], b = beta1[col];
  float cs = 0.f, am = 0.f;
#pragma unroll 4
  for (int i = 0; i < MD_RPT; ++i) {
    const int r = rg * MD_RPT + i;
    if (r < cnt) {
      const int64_t o = (r0 + r) * MD_H + col;
      const float xhat = (y1[o] - mu) * rs;
      float d = da1[o];
      if (mask) d = mask[o] ? d * drop_scale : 0.f;
      if (!(xhat * g + b > 0.f)) d = 0.f;
      const float v = rs * g * (d - S1 * invB - xhat * (S2 * invB));
      dy1[o] = v;
      cs += v;
      am = fmaxf(am, fabsf(v));
    }
  }
  s_scr[rg * MD_H + col] = cs;
  __syncthreads();
  if (rg == 0)
    ppart[(int64_t)blockIdx.x * prec + dbe_off + col] =
        (s_scr[col] + s_scr[MD_H + col]) + (s_scr[2 * MD_H + col] + s_scr[3 * MD_H + col]);
  if (absmax_part) {  // max |dy1| of this tile (operand scale of the tensor-core dW kernel downstream)
    __syncthreads();
    am = warp_max(am);
    if ((threadIdx.x & 31) == 0) s_scr[threadIdx.x >> 5] = am;
    __syncthreads();
    if (threadIdx.x < 32) {
      am = threadIdx.x < MD_THREADS / 32 ? s_scr[threadIdx.x] : 0.f;
      am = warp_max(am);
      if (threadIdx.x == 0) absmax_part[blockIdx.x] = am;
    }
  }
}

}  // namespace scvi

extern "C" {

int64_t scvi_mid_n_tiles(int64_t B) { return scvi::cdiv(B, scvi::MD_ROWS); }

int64_t scvi_mid_ws_bytes(int64_t B, int64_t L, int64_t ncols) {
  const int64_t T = scvi::cdiv(B, scvi::MD_ROWS);
  const int64_t prec = scvi::md_rec_floats((int)ncols, (int)L);
  return 4 * (3 * T * 2 * scvi::MD_H + B * scvi::MD_H + (T + 1) * prec) + 1024;
}

int scvi_mid_fwd(const float* y1, const float* bn_partials1, int64_t B, int64_t H, int64_t L, const float* gamma1,
                 const float* beta1, float* running_mean1, float* running_var1, const uint8_t* drop_mask,
                 float drop_scale, const float* noise, const float* Wm, const float* bm, const float* Wv,
                 const float* bv, float var_eps, const float* Wd, int64_t n_w_cols, const float* bd,
                 const int32_t* cov_cols, const float* gamma2, const float* beta2, float* running_mean2,
                 float* running_var2, float bn_eps, float momentum, float* a1, float* q_m, float* q_v, float* z,
                 float* kl, float* y2, float* h, float* mean1, float* rstd1, float* mean2, float* rstd2, void* ws,
                 void* stream) {
  using namespace scvi;
  SCVI_REQUIRE(y1 && bn_partials1 && gamma1 && beta1 && noise && Wm && bm && Wv && bv && Wd && gamma2 && beta2 && a1 &&
                   q_m && q_v && z && kl && y2 && h && mean1 && rstd1 && mean2 && rstd2 && ws,
               "scvi_mid_fwd: NULL pointer");
  if (H != MD_H || L < 1 || L > MD_LMAX || n_w_cols < L || n_w_cols >= MD_WD_LD) {
    set_error("scvi_mid_fwd: needs n_hidden == 128, n_latent <= %d, n_latent + n_batch <= %d", MD_LMAX, MD_WD_LD - 1);
    return SCVI_EUNSUPPORTED;
  }
  SCVI_REQUIRE(B > 0, "scvi_mid_fwd: empty batch");
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t T = cdiv(B, MD_ROWS);
  float* part2 = (float*)ws;
  MidFwdArgs A;
  A.y1 = y1; A.part1 = bn_partials1; A.gamma1 = gamma1; A.beta1 = beta1; A.rm1 = running_mean1; A.rv1 = running_var1;
  A.mask = drop_mask; A.drop_scale = drop_scale; A.noise = noise; A.Wm = Wm; A.bm = bm; A.Wv = Wv; A.bv = bv;
  A.var_eps = var_eps; A.Wd = Wd; A.bd = bd; A.cov_cols = cov_cols; A.ncols = (int)n_w_cols; A.bn_eps = bn_eps;
  A.momentum = momentum; A.B = B; A.L = (int)L; A.a1 = a1; A.q_m = q_m; A.q_v = q_v; A.z = z; A.kl = kl; A.y2 = y2;
  A.part2 = part2; A.mean1 = mean1; A.rstd1 = rstd1;
  const size_t smem = sizeof(float) * ((size_t)MD_ROWS * MD_H + 2 * MD_LMAX * 129 + MD_H * MD_WD_LD +
                                       MD_ROWS * 2 * MD_LMAX + 2 * MD_ROWS * MD_LMAX + 12 * MD_H);
  static std::atomic<uint64_t> optin{0};
  if (int rc = smem_optin((const void*)mid_fwd_kernel, (int)smem, optin, "scvi_mid_fwd")) return rc;
  mid_fwd_kernel<<<(unsigned)(md_clusters(B) * MD_CL), MD_THREADS, smem, st>>>(A);
  SCVI_LAUNCH_CHECK("scvi_mid_fwd(stage 1)");
  mid_fwd2_kernel<<<(unsigned)T, MD_THREADS, 0, st>>>(y2, part2, gamma2, beta2, running_mean2, running_var2, bn_eps,
                                                     momentum, B, h, mean2, rstd2);
  SCVI_LAUNCH_CHECK("scvi_mid_fwd(stage 2)");
  return SCVI_OK;
}

int scvi_mid_bwd(const float* dh, const float* y2, const float* mean2, const float* rstd2, const float* gamma2,
                 const float* beta2, const float* z, const float* Wd, int64_t n_w_cols, const int32_t* cov_cols,
                 const float* g_kl, const float* q_m, const float* q_v, const float* noise, float var_eps,
                 const float* Wm, const float* Wv, const float* a1, const float* y1, const float* mean1,
                 const float* rstd1, const float* gamma1, const float* beta1, const uint8_t* drop_mask,
                 float drop_scale, int64_t B, int64_t H, int64_t L, float* dy1, float* dgamma2, float* dbeta2,
                 float* dWd, float* dbd, float* dWm, float* dbm, float* dWv, float* dbv, float* dgamma1, float* dbeta1,
                 float* db_enc, float* dy1_absmax_part, void* ws, void* stream) {
  using namespace scvi;
  SCVI_REQUIRE(dh && y2 && mean2 && rstd2 && gamma2 && beta2 && z && Wd && q_m && q_v && noise && Wm && Wv && a1 && y1 &&
                   mean1 && rstd1 && gamma1 && beta1 && dy1 && ws,
               "scvi_mid_bwd: NULL pointer");
  if (H != MD_H || L < 1 || L > MD_LMAX || n_w_cols < L || n_w_cols >= MD_WD_LD) {
    set_error("scvi_mid_bwd: needs n_hidden == 128, n_latent <= %d, n_latent + n_batch <= %d", MD_LMAX, MD_WD_LD - 1);
    return SCVI_EUNSUPPORTED;
  }
  SCVI_REQUIRE(B > 0, "scvi_mid_bwd: empty batch");
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t T = cdiv(B, MD_ROWS);
  const int ncols = (int)n_w_cols;
  const int64_t prec = md_rec_floats(ncols, (int)L);
  float* w = (float*)ws;
  float* bpart2 = w + T * 2 * MD_H;  // [0, T*2*128) is part2 of the forward
  float* bpart1 = bpart2 + T * 2 * MD_H;
  float* da1 = bpart1 + T * 2 * MD_H;
  float* ppart = da1 + B * MD_H;
  mid_bwd0_kernel<<<(unsigned)(md_clusters(B) * MD_CL), MD_THREADS, 0, st>>>(dh, y2, mean2, rstd2, gamma2, beta2, B, bpart2);
  SCVI_LAUNCH_CHECK("scvi_mid_bwd(stage 0)");
  MidBwdArgs A;
  A.dh = dh; A.y2 = y2; A.mean2 = mean2; A.rstd2 = rstd2; A.gamma2 = gamma2; A.beta2 = beta2; A.bpart2 = bpart2;
  A.z = z; A.Wd = Wd; A.cov_cols = cov_cols; A.ncols = ncols; A.g_kl = g_kl; A.q_m = q_m; A.q_v = q_v; A.noise = noise;
  A.var_eps = var_eps; A.Wm = Wm; A.Wv = Wv; A.a1 = a1; A.y1 = y1; A.mean1 = mean1; A.rstd1 = rstd1; A.gamma1 = gamma1;
  A.beta1 = beta1; A.mask = drop_mask; A.drop_scale = drop_scale; A.B = B; A.L = (int)L; A.da1 = da1; A.bpart1 = bpart1;
  A.ppart = ppart; A.prec = prec; A.dgamma2 = dgamma2; A.dbeta2 = dbeta2;
  const size_t smem = sizeof(float) * ((size_t)MD_ROWS * MD_H + 2 * MD_LMAX * 129 + MD_H * MD_WD_LD + MD_ROWS * MD_LMAX +
                                       MD_ROWS * 2 * MD_LMAX + 12 * MD_H);
  static std::atomic<uint64_t> optin{0};
  if (int rc = smem_optin((const void*)mid_bwd1_kernel, (int)smem, optin, "scvi_mid_bwd")) return rc;
  mid_bwd1_kernel<<<(unsigned)(md_clusters(B) * MD_CL), MD_THREADS, smem, st>>>(A);
  SCVI_LAUNCH_CHECK("scvi_mid_bwd(stage 1)");
  const int64_t dbe_off = prec - MD_H;
  mid_bwd2_kernel<<<(unsigned)T, MD_THREADS, 0, st>>>(da1, y1, mean1, rstd1, gamma1, beta1, drop_mask, drop_scale,
                                                     bpart1, B, dy1, dgamma1, dbeta1, ppart, prec, dbe_off,
                                                     dy1_absmax_part);
  SCVI_LAUNCH_CHECK("scvi_mid_bwd(stage 2)");
  // one deterministic column sum over the per-tile records, scattered to the 7 gradient tensors
  ColsumSegs segs{};
  int64_t o = 0;
  auto add = [&](float* dst, int64_t n) {
    segs.dst[segs.n] = dst; segs.begin[segs.n] = o; segs.end[segs.n] = o + n; ++segs.n;
    o += n;
  };
  add(dWd, (int64_t)MD_H * ncols);
  add(dbd, MD_H);
  add(dWm, L * MD_H);
  add(dWv, L * MD_H);
  add(dbm, L);
  add(dbv, L);
  add(db_enc, MD_H);
  return launch_colsum(ppart, T, prec, prec, segs, st, "scvi_mid_bwd(reduce)");
}
}
