// Fused decoder heads + gene softmax + library scaling + NB/ZINB negative log-likelihood.
// fp32 FFMA tiles (SCVI_PREC_FP32, the parity path).  Replaces DecoderSCVI.forward's heads
// (reference src/scvi/nn/_base_components.py:436-441), `px_r = exp(self.px_r)`
// (src/scvi/module/_vae.py:516-518), log_zinb_positive / log_nb_positive
// (src/scvi/distributions/_negative_binomial.py:36-135) and `-log_prob(x).sum(-1)`
// (_vae.py:571).  No [B,G] tensor is written to HBM: the logits are recomputed per pass.
//
//   forward  : CTA = 64 rows.  pass 1: scale-head GEMM -> online row max / sum-exp.
//              pass 2: both head GEMMs -> mu = exp(s - max) * exp(library)/sumexp -> ll,
//              row sums of ll and of S = sum_g mu * d(reconst)/d(mu) (softmax backward term).
//   backward : A (CTA = 64 rows, all genes)  -> dh = ds Ws + dp Wp        (no atomics)
//              B (CTA = 64 genes, row split) -> dWs, dWp, dbs, dbp, dpx_r (partial slabs,
//              reduced in fixed order)
#include "decoder_math.cuh"

namespace scvi {

constexpr int DT = 64;         // tile edge: rows per CTA (fwd / bwd A), genes per tile
constexpr int DTHREADS = 256;  // 16 x 16 threads, 4 x 4 strided micro-tile each

// 64x64 logits tile: acc[i][j] += sum_k hs[ty+16i][k] * ws[tx+16j][k]
template <bool TWO>
__device__ __forceinline__ void tile_gemm(const float* __restrict__ hs, const float* __restrict__ ws,
                                          const float* __restrict__ wp, int H, int ldh, int tx, int ty,
                                          float (&as)[4][4], float (&ap)[4][4]) {
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      as[i][j] = 0.f;
      if (TWO) ap[i][j] = 0.f;
    }
#pragma unroll 2
  for (int k = 0; k < H; k += 4) {
    float4 a[4], b[4], c[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) a[i] = *reinterpret_cast<const float4*>(&hs[(ty + 16 * i) * ldh + k]);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      b[j] = *reinterpret_cast<const float4*>(&ws[(tx + 16 * j) * ldh + k]);
      if (TWO) c[j] = *reinterpret_cast<const float4*>(&wp[(tx + 16 * j) * ldh + k]);
    }
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        as[i][j] = fmaf(a[i].x, b[j].x, as[i][j]);
        as[i][j] = fmaf(a[i].y, b[j].y, as[i][j]);
        as[i][j] = fmaf(a[i].z, b[j].z, as[i][j]);
        as[i][j] = fmaf(a[i].w, b[j].w, as[i][j]);
        if (TWO) {
          ap[i][j] = fmaf(a[i].x, c[j].x, ap[i][j]);
          ap[i][j] = fmaf(a[i].y, c[j].y, ap[i][j]);
          ap[i][j] = fmaf(a[i].z, c[j].z, ap[i][j]);
          ap[i][j] = fmaf(a[i].w, c[j].w, ap[i][j]);
        }
      }
  }
}

// copy `nrows` rows of H floats (row stride ld_src) into smem rows of stride ldh; rows >= valid are zeroed
__device__ __forceinline__ void load_rows(float* __restrict__ dst, int ldh, const float* __restrict__ src,
                                          int64_t ld_src, int H, int valid, int tid) {
  const int h4 = H >> 2;
  for (int e = tid; e < DT * h4; e += DTHREADS) {
    const int r = e / h4, q = e % h4;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (r < valid) v = *reinterpret_cast<const float4*>(src + (int64_t)r * ld_src + 4 * q);
    *reinterpret_cast<float4*>(&dst[r * ldh + 4 * q]) = v;
  }
}

// ---------------------------------------------------------------------------------------
// forward
// ---------------------------------------------------------------------------------------
template <bool ZINB>
__global__ void __launch_bounds__(DTHREADS)
    decoder_nll_fwd_kernel(const float* __restrict__ h, int64_t ldh_g, int64_t B, int H, int64_t G,
                           const float* __restrict__ Ws, const float* __restrict__ bs,
                           const float* __restrict__ Wp, const float* __restrict__ bp,
                           const float* __restrict__ px_r, const float* __restrict__ x, int64_t ldx,
                           const int64_t* __restrict__ row_idx, const float* __restrict__ library, float eps,
                           float* __restrict__ reconst, float* __restrict__ row_stats) {
  extern __shared__ __align__(16) float smem[];
  const int ldh = H + 4;
  float* hs = smem;                 // [DT][ldh]
  float* ws = hs + DT * ldh;        // [DT][ldh]
  float* wp = ws + DT * ldh;        // [DT][ldh] (ZINB only)
  float* gcs = ZINB ? wp + DT * ldh : ws + DT * ldh;  // [DT][5]: theta, log_theta, lg_theta, bs, bp
  __shared__ int64_t rows[DT];

  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int64_t r0 = (int64_t)blockIdx.x * DT;
  const int valid = (int)min((int64_t)DT, B - r0);
  if (tid < DT) rows[tid] = (tid < valid) ? (row_idx ? row_idx[r0 + tid] : r0 + tid) : -1;
  load_rows(hs, ldh, h + r0 * ldh_g, ldh_g, H, valid, tid);

  float as[4][4], ap[4][4];
  // ---- pass 1: online log-sum-exp of the scale logits
  float m[4], l[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    m[i] = -INFINITY;
    l[i] = 0.f;
  }
  for (int64_t g0 = 0; g0 < G; g0 += DT) {
    __syncthreads();
    load_rows(ws, ldh, Ws + g0 * H, H, H, (int)min((int64_t)DT, G - g0), tid);
    if (tid < DT) gcs[tid * 5 + 3] = (g0 + tid < G) ? bs[g0 + tid] : 0.f;
    __syncthreads();
    tile_gemm<false>(hs, ws, nullptr, H, ldh, tx, ty, as, ap);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float tm = -INFINITY;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        as[i][j] += gcs[(tx + 16 * j) * 5 + 3];
        if (g0 + tx + 16 * j < G) tm = fmaxf(tm, as[i][j]);
      }
      if (tm > m[i]) {
        l[i] *= expf(m[i] - tm);  // m = -inf first time: exp(-inf) = 0
        m[i] = tm;
      }
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (g0 + tx + 16 * j < G) l[i] += expf(as[i][j] - m[i]);
    }
  }
  // merge (m, l) over the 16 lanes that share a row
  RowC rc[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) {
      const float mo = __shfl_xor_sync(0xffffffffu, m[i], o);
      const float lo = __shfl_xor_sync(0xffffffffu, l[i], o);
      const float mn = fmaxf(m[i], mo);
      l[i] = (m[i] == -INFINITY ? 0.f : l[i] * expf(m[i] - mn)) + (mo == -INFINITY ? 0.f : lo * expf(mo - mn));
      m[i] = mn;
    }
    const int rr = ty + 16 * i;
    const float lib = (rr < valid) ? library[r0 + rr] : 0.f;
    rc[i].m = m[i];
    rc[i].invl = 1.f / l[i];
    rc[i].c = expf(lib) / l[i];  // exp(library) * softmax = exp(s - m) * c
  }

  // ---- pass 2: likelihood
  float nll[4] = {0.f, 0.f, 0.f, 0.f}, S[4] = {0.f, 0.f, 0.f, 0.f};
  for (int64_t g0 = 0; g0 < G; g0 += DT) {
    __syncthreads();
    const int gvalid = (int)min((int64_t)DT, G - g0);
    load_rows(ws, ldh, Ws + g0 * H, H, H, gvalid, tid);
    if (ZINB) load_rows(wp, ldh, Wp + g0 * H, H, H, gvalid, tid);
    if (tid < DT) {
      const bool ok = tid < gvalid;
      const GeneConst gc = make_gene_const(ok ? expf(px_r[g0 + tid]) : 1.f, eps);
      gcs[tid * 5 + 0] = gc.theta;
      gcs[tid * 5 + 1] = gc.log_theta;
      gcs[tid * 5 + 2] = gc.lg_theta;
      gcs[tid * 5 + 3] = ok ? bs[g0 + tid] : 0.f;
      gcs[tid * 5 + 4] = (ZINB && ok) ? bp[g0 + tid] : 0.f;
    }
    __syncthreads();
    tile_gemm<ZINB>(hs, ws, wp, H, ldh, tx, ty, as, ap);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int gl = tx + 16 * j;
      if (gl >= gvalid) continue;
      GeneConst gc;
      gc.theta = gcs[gl * 5 + 0];
      gc.log_theta = gcs[gl * 5 + 1];
      gc.lg_theta = gcs[gl * 5 + 2];
      const float bsv = gcs[gl * 5 + 3], bpv = gcs[gl * 5 + 4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int rr = ty + 16 * i;
        if (rr >= valid) continue;
        const float xv = ld_stream(x + rows[rr] * ldx + g0 + gl);
        const float mu = expf(as[i][j] + bsv - rc[i].m) * rc[i].c;
        float ll, dmu;
        elem_fwd<ZINB>(xv, mu, ZINB ? ap[i][j] + bpv : 0.f, gc, eps, ll, dmu);
        nll[i] -= ll;
        S[i] -= mu * dmu;  // d(reconst)/d(mu) = -d(ll)/d(mu)
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) {
      nll[i] += __shfl_xor_sync(0xffffffffu, nll[i], o);
      S[i] += __shfl_xor_sync(0xffffffffu, S[i], o);
    }
    const int rr = ty + 16 * i;
    if (tx == 0 && rr < valid) {
      reconst[r0 + rr] = nll[i];
      float4 st = make_float4(rc[i].m, l[i], rc[i].c, S[i]);
      *reinterpret_cast<float4*>(row_stats + (r0 + rr) * 4) = st;
    }
  }
}

// ---------------------------------------------------------------------------------------
// backward A: dh[64 rows, H] = sum over all genes of ds * Ws + dp * Wp
// ---------------------------------------------------------------------------------------
template <bool ZINB, int NH4 /* H / 64 */>
__global__ void __launch_bounds__(DTHREADS)
    decoder_nll_bwd_dh_kernel(const float* __restrict__ h, int64_t ldh_g, int64_t B, int64_t G,
                              const float* __restrict__ Ws, const float* __restrict__ bs,
                              const float* __restrict__ Wp, const float* __restrict__ bp,
                              const float* __restrict__ px_r, const float* __restrict__ x, int64_t ldx,
                              const int64_t* __restrict__ row_idx, const float* __restrict__ row_stats,
                              const float* __restrict__ g_reconst, float eps, float* __restrict__ dh,
                              int64_t lddh) {
  constexpr int H = 64 * NH4;
  constexpr int ldh = H + 4;
  constexpr int ldd = DT + 4;
  extern __shared__ __align__(16) float smem[];
  float* hs = smem;
  float* ws = hs + DT * ldh;
  float* wp = ws + DT * ldh;
  float* dss = ZINB ? wp + DT * ldh : ws + DT * ldh;  // [DT rows][ldd]
  float* dps = dss + DT * ldd;
  float* gcs = ZINB ? dps + DT * ldd : dss + DT * ldd;  // [DT][5]
  __shared__ int64_t rows[DT];

  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int64_t r0 = (int64_t)blockIdx.x * DT;
  const int valid = (int)min((int64_t)DT, B - r0);
  if (tid < DT) rows[tid] = (tid < valid) ? (row_idx ? row_idx[r0 + tid] : r0 + tid) : -1;
  load_rows(hs, ldh, h + r0 * ldh_g, ldh_g, H, valid, tid);

  RowC rc[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int rr = ty + 16 * i;
    if (rr < valid) {
      const float4 st = *reinterpret_cast<const float4*>(row_stats + (r0 + rr) * 4);
      rc[i].m = st.x;
      rc[i].invl = 1.f / st.y;
      rc[i].c = st.z;
      rc[i].S = st.w;
      rc[i].u = g_reconst ? g_reconst[r0 + rr] : 1.f;
    } else {
      rc[i].m = 0.f; rc[i].invl = 0.f; rc[i].c = 0.f; rc[i].S = 0.f; rc[i].u = 0.f;
    }
  }
  float acc[4][4 * NH4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int q = 0; q < 4 * NH4; ++q) acc[i][q] = 0.f;

  float as[4][4], ap[4][4];
  for (int64_t g0 = 0; g0 < G; g0 += DT) {
    __syncthreads();
    const int gvalid = (int)min((int64_t)DT, G - g0);
    load_rows(ws, ldh, Ws + g0 * H, H, H, gvalid, tid);
    if (ZINB) load_rows(wp, ldh, Wp + g0 * H, H, H, gvalid, tid);
    if (tid < DT) {
      const bool ok = tid < gvalid;
      const GeneConst gc = make_gene_const(ok ? expf(px_r[g0 + tid]) : 1.f, eps);
      gcs[tid * 5 + 0] = gc.theta;
      gcs[tid * 5 + 1] = gc.log_theta;
      gcs[tid * 5 + 2] = gc.lg_theta;
      gcs[tid * 5 + 3] = ok ? bs[g0 + tid] : 0.f;
      gcs[tid * 5 + 4] = (ZINB && ok) ? bp[g0 + tid] : 0.f;
    }
    __syncthreads();
    tile_gemm<ZINB>(hs, ws, wp, H, ldh, tx, ty, as, ap);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int gl = tx + 16 * j;
      GeneConst gc;
      gc.theta = gcs[gl * 5 + 0];
      gc.log_theta = gcs[gl * 5 + 1];
      gc.lg_theta = gcs[gl * 5 + 2];
      const float bsv = gcs[gl * 5 + 3], bpv = gcs[gl * 5 + 4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int rr = ty + 16 * i;
        float ds = 0.f, dp = 0.f, dr;
        if (gl < gvalid && rr < valid) {
          const float xv = ld_stream(x + rows[rr] * ldx + g0 + gl);
          elem_bwd<ZINB>(xv, as[i][j] + bsv, ZINB ? ap[i][j] + bpv : 0.f, rc[i], gc, eps, ds, dp, dr);
        }
        dss[rr * ldd + gl] = ds;
        if (ZINB) dps[rr * ldd + gl] = dp;
      }
    }
    __syncthreads();
    // dh[rows ty+16i][cols 64*q + 4*tx .. +3] += sum_n ds[row][n] * ws[n][col]
#pragma unroll 2
    for (int n = 0; n < DT; ++n) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        a[i] = dss[(ty + 16 * i) * ldd + n];
        if (ZINB) b[i] = dps[(ty + 16 * i) * ldd + n];
      }
#pragma unroll
      for (int q = 0; q < NH4; ++q) {
        const float4 w = *reinterpret_cast<const float4*>(&ws[n * ldh + 64 * q + 4 * tx]);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          acc[i][4 * q + 0] = fmaf(a[i], w.x, acc[i][4 * q + 0]);
          acc[i][4 * q + 1] = fmaf(a[i], w.y, acc[i][4 * q + 1]);
          acc[i][4 * q + 2] = fmaf(a[i], w.z, acc[i][4 * q + 2]);
          acc[i][4 * q + 3] = fmaf(a[i], w.w, acc[i][4 * q + 3]);
        }
        if (ZINB) {
          const float4 v = *reinterpret_cast<const float4*>(&wp[n * ldh + 64 * q + 4 * tx]);
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            acc[i][4 * q + 0] = fmaf(b[i], v.x, acc[i][4 * q + 0]);
            acc[i][4 * q + 1] = fmaf(b[i], v.y, acc[i][4 * q + 1]);
            acc[i][4 * q + 2] = fmaf(b[i], v.z, acc[i][4 * q + 2]);
            acc[i][4 * q + 3] = fmaf(b[i], v.w, acc[i][4 * q + 3]);
          }
        }
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int rr = ty + 16 * i;
    if (rr >= valid) continue;
#pragma unroll
    for (int q = 0; q < NH4; ++q)
      *reinterpret_cast<float4*>(dh + (r0 + rr) * lddh + 64 * q + 4 * tx) =
          make_float4(acc[i][4 * q], acc[i][4 * q + 1], acc[i][4 * q + 2], acc[i][4 * q + 3]);
  }
}

// ---------------------------------------------------------------------------------------
// backward B: CTA owns 64 genes (weights resident), loops over its share of the rows.
//   dWs[g, :] = sum_rows ds[row, g] * h[row, :]   (same for dWp), dbs, dbp, dpx_r column sums
// ---------------------------------------------------------------------------------------
template <bool ZINB, int NH4>
__global__ void __launch_bounds__(DTHREADS)
    decoder_nll_bwd_dw_kernel(const float* __restrict__ h, int64_t ldh_g, int64_t B, int64_t G,
                              const float* __restrict__ Ws, const float* __restrict__ bs,
                              const float* __restrict__ Wp, const float* __restrict__ bp,
                              const float* __restrict__ px_r, const float* __restrict__ x, int64_t ldx,
                              const int64_t* __restrict__ row_idx, const float* __restrict__ row_stats,
                              const float* __restrict__ g_reconst, float eps, int64_t rows_per_split,
                              float* __restrict__ slabW /*[splits][nheads][G][H]*/,
                              float* __restrict__ slabv /*[splits][3][G]: dbs, dbp, dpx_r*/) {
  constexpr int H = 64 * NH4;
  constexpr int ldh = H + 4;
  constexpr int ldd = DT + 4;
  extern __shared__ __align__(16) float smem[];
  float* hs = smem;
  float* ws = hs + DT * ldh;
  float* wp = ws + DT * ldh;
  float* dss = ZINB ? wp + DT * ldh : ws + DT * ldh;
  float* dps = dss + DT * ldd;
  float* rcs = ZINB ? dps + DT * ldd : dss + DT * ldd;  // [DT][5] row constants
  __shared__ int64_t rows[DT];

  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int64_t g0 = (int64_t)blockIdx.x * DT;
  const int gvalid = (int)min((int64_t)DT, G - g0);
  const int64_t b_begin = (int64_t)blockIdx.y * rows_per_split;
  const int64_t b_end = min(B, b_begin + rows_per_split);
  const int nheads = ZINB ? 2 : 1;

  load_rows(ws, ldh, Ws + g0 * H, H, H, gvalid, tid);
  if (ZINB) load_rows(wp, ldh, Wp + g0 * H, H, H, gvalid, tid);
  GeneConst gc[4];
  float bsv[4], bpv[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int gl = tx + 16 * j;
    const bool ok = gl < gvalid;
    gc[j] = make_gene_const(ok ? expf(px_r[g0 + gl]) : 1.f, eps);
    bsv[j] = ok ? bs[g0 + gl] : 0.f;
    bpv[j] = (ZINB && ok) ? bp[g0 + gl] : 0.f;
  }
  float accS[4][4 * NH4], accP[4][4 * NH4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int q = 0; q < 4 * NH4; ++q) {
      accS[i][q] = 0.f;
      if (ZINB) accP[i][q] = 0.f;
    }
  float cs[4] = {0.f, 0.f, 0.f, 0.f}, cp[4] = {0.f, 0.f, 0.f, 0.f}, cr[4] = {0.f, 0.f, 0.f, 0.f};

  float as[4][4], ap[4][4];
  for (int64_t rb = b_begin; rb < b_end; rb += DT) {
    __syncthreads();
    const int valid = (int)min((int64_t)DT, b_end - rb);
    load_rows(hs, ldh, h + rb * ldh_g, ldh_g, H, valid, tid);
    if (tid < DT) {
      const bool ok = tid < valid;
      rows[tid] = ok ? (row_idx ? row_idx[rb + tid] : rb + tid) : -1;
      float4 st = ok ? *reinterpret_cast<const float4*>(row_stats + (rb + tid) * 4) : make_float4(0.f, 1.f, 0.f, 0.f);
      rcs[tid * 5 + 0] = st.x;
      rcs[tid * 5 + 1] = 1.f / st.y;
      rcs[tid * 5 + 2] = st.z;
      rcs[tid * 5 + 3] = st.w;
      rcs[tid * 5 + 4] = ok ? (g_reconst ? g_reconst[rb + tid] : 1.f) : 0.f;
    }
    __syncthreads();
    tile_gemm<ZINB>(hs, ws, wp, H, ldh, tx, ty, as, ap);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int rr = ty + 16 * i;
      RowC rc;
      rc.m = rcs[rr * 5 + 0];
      rc.invl = rcs[rr * 5 + 1];
      rc.c = rcs[rr * 5 + 2];
      rc.S = rcs[rr * 5 + 3];
      rc.u = rcs[rr * 5 + 4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int gl = tx + 16 * j;
        float ds = 0.f, dp = 0.f, dr = 0.f;
        if (gl < gvalid && rr < valid) {
          const float xv = ld_stream(x + rows[rr] * ldx + g0 + gl);
          elem_bwd<ZINB>(xv, as[i][j] + bsv[j], ZINB ? ap[i][j] + bpv[j] : 0.f, rc, gc[j], eps, ds, dp, dr);
        }
        dss[rr * ldd + gl] = ds;
        if (ZINB) dps[rr * ldd + gl] = dp;
        cs[j] += ds;
        cp[j] += dp;
        cr[j] += dr;
      }
    }
    __syncthreads();
    // dW[genes ty+16i][cols 64q + 4tx..] += sum_r ds[r][gene] * hs[r][col]
#pragma unroll 2
    for (int r = 0; r < DT; ++r) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        a[i] = dss[r * ldd + ty + 16 * i];
        if (ZINB) b[i] = dps[r * ldd + ty + 16 * i];
      }
#pragma unroll
      for (int q = 0; q < NH4; ++q) {
        const float4 hv = *reinterpret_cast<const float4*>(&hs[r * ldh + 64 * q + 4 * tx]);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          accS[i][4 * q + 0] = fmaf(a[i], hv.x, accS[i][4 * q + 0]);
          accS[i][4 * q + 1] = fmaf(a[i], hv.y, accS[i][4 * q + 1]);
          accS[i][4 * q + 2] = fmaf(a[i], hv.z, accS[i][4 * q + 2]);
          accS[i][4 * q + 3] = fmaf(a[i], hv.w, accS[i][4 * q + 3]);
          if (ZINB) {
            accP[i][4 * q + 0] = fmaf(b[i], hv.x, accP[i][4 * q + 0]);
            accP[i][4 * q + 1] = fmaf(b[i], hv.y, accP[i][4 * q + 1]);
            accP[i][4 * q + 2] = fmaf(b[i], hv.z, accP[i][4 * q + 2]);
            accP[i][4 * q + 3] = fmaf(b[i], hv.w, accP[i][4 * q + 3]);
          }
        }
      }
    }
  }
  // write the dW partial slab of this split
  float* oS = slabW + ((int64_t)blockIdx.y * nheads + 0) * G * H;
  float* oP = slabW + ((int64_t)blockIdx.y * nheads + 1) * G * H;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int gl = ty + 16 * i;
    if (gl >= gvalid) continue;
#pragma unroll
    for (int q = 0; q < NH4; ++q) {
      *reinterpret_cast<float4*>(oS + (g0 + gl) * H + 64 * q + 4 * tx) =
          make_float4(accS[i][4 * q], accS[i][4 * q + 1], accS[i][4 * q + 2], accS[i][4 * q + 3]);
      if (ZINB)
        *reinterpret_cast<float4*>(oP + (g0 + gl) * H + 64 * q + 4 * tx) =
            make_float4(accP[i][4 * q], accP[i][4 * q + 1], accP[i][4 * q + 2], accP[i][4 * q + 3]);
    }
  }
  // column sums: reduce the 16 ty-threads sharing a gene through smem (reuse dss)
  __syncthreads();
  float* red = dss;  // [3][16 ty][64 genes]
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    red[(0 * 16 + ty) * DT + tx + 16 * j] = cs[j];
    red[(1 * 16 + ty) * DT + tx + 16 * j] = cp[j];
    red[(2 * 16 + ty) * DT + tx + 16 * j] = cr[j];
  }
  __syncthreads();
  if (tid < 3 * DT) {
    const int which = tid / DT, gl = tid % DT;
    if (gl < gvalid) {
      float s = 0.f;
#pragma unroll
      for (int t = 0; t < 16; ++t) s += red[(which * 16 + t) * DT + gl];
      slabv[((int64_t)blockIdx.y * 3 + which) * G + g0 + gl] = s;
    }
  }
}

// sum the split slabs in fixed order
__global__ void decoder_bwd_reduce_kernel(const float* __restrict__ slabW, const float* __restrict__ slabv,
                                          int64_t splits, int nheads, int64_t GH, int64_t G,
                                          float* __restrict__ dWs, float* __restrict__ dWp,
                                          float* __restrict__ dbs, float* __restrict__ dbp,
                                          float* __restrict__ dpx_r) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e < GH) {
    float s = 0.f, p = 0.f;
    for (int64_t t = 0; t < splits; ++t) {
      s += slabW[(t * nheads + 0) * GH + e];
      if (nheads == 2) p += slabW[(t * nheads + 1) * GH + e];
    }
    if (dWs) dWs[e] = s;
    if (nheads == 2 && dWp) dWp[e] = p;
  }
  if (e < G) {
    float a = 0.f, b = 0.f, c = 0.f;
    for (int64_t t = 0; t < splits; ++t) {
      a += slabv[(t * 3 + 0) * G + e];
      b += slabv[(t * 3 + 1) * G + e];
      c += slabv[(t * 3 + 2) * G + e];
    }
    if (dbs) dbs[e] = a;
    if (nheads == 2 && dbp) dbp[e] = b;
    if (dpx_r) dpx_r[e] = c;
  }
}

// ---------------------------------------------------------------------------------------
// materialising variant for inference-time getters
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(DTHREADS)
    decoder_params_kernel(const float* __restrict__ h, int64_t ldh_g, int64_t B, int H, int64_t G,
                          const float* __restrict__ Ws, const float* __restrict__ bs,
                          const float* __restrict__ Wp, const float* __restrict__ bp,
                          const float* __restrict__ library, float* __restrict__ px_scale,
                          float* __restrict__ px_rate, float* __restrict__ px_dropout, int64_t ldo) {
  extern __shared__ __align__(16) float smem[];
  const int ldh = H + 4;
  float* hs = smem;
  float* ws = hs + DT * ldh;
  float* bsm = ws + DT * ldh;  // [DT]
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int64_t r0 = (int64_t)blockIdx.x * DT;
  const int valid = (int)min((int64_t)DT, B - r0);
  load_rows(hs, ldh, h + r0 * ldh_g, ldh_g, H, valid, tid);
  float as[4][4], ap[4][4];
  float m[4], l[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    m[i] = -INFINITY;
    l[i] = 0.f;
  }
  const bool need_softmax = px_scale || px_rate;
  for (int pass = need_softmax ? 0 : 2; pass < 3; ++pass) {
    // pass 0: LSE of the scale head; pass 1: write scale / rate; pass 2: dropout head
    if (pass == 2 && !px_dropout) break;
    const float* Wsel = (pass == 2) ? Wp : Ws;
    const float* bsel = (pass == 2) ? bp : bs;
    for (int64_t g0 = 0; g0 < G; g0 += DT) {
      __syncthreads();
      const int gvalid = (int)min((int64_t)DT, G - g0);
      load_rows(ws, ldh, Wsel + g0 * H, H, H, gvalid, tid);
      if (tid < DT) bsm[tid] = (tid < gvalid) ? bsel[g0 + tid] : 0.f;
      __syncthreads();
      tile_gemm<false>(hs, ws, nullptr, H, ldh, tx, ty, as, ap);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int rr = ty + 16 * i;
        if (pass == 0) {
          float tm = -INFINITY;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            as[i][j] += bsm[tx + 16 * j];
            if (tx + 16 * j < gvalid) tm = fmaxf(tm, as[i][j]);
          }
          if (tm > m[i]) {
            l[i] *= expf(m[i] - tm);
            m[i] = tm;
          }
#pragma unroll
          for (int j = 0; j < 4; ++j)
            if (tx + 16 * j < gvalid) l[i] += expf(as[i][j] - m[i]);
        } else if (rr < valid) {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int gl = tx + 16 * j;
            if (gl >= gvalid) continue;
            const float v = as[i][j] + bsm[gl];
            const int64_t o = (r0 + rr) * ldo + g0 + gl;
            if (pass == 1) {
              const float sc = expf(v - m[i]) / l[i];
              if (px_scale) px_scale[o] = sc;
              if (px_rate) px_rate[o] = expf(library[r0 + rr]) * sc;
            } else {
              px_dropout[o] = v;
            }
          }
        }
      }
    }
    if (pass == 0) {
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) {
          const float mo = __shfl_xor_sync(0xffffffffu, m[i], o);
          const float lo = __shfl_xor_sync(0xffffffffu, l[i], o);
          const float mn = fmaxf(m[i], mo);
          l[i] = (m[i] == -INFINITY ? 0.f : l[i] * expf(m[i] - mn)) + (mo == -INFINITY ? 0.f : lo * expf(mo - mn));
          m[i] = mn;
        }
    }
  }
}

static int64_t dw_row_splits(int64_t B, int64_t G) {
  const int64_t tiles = cdiv(G, DT);
  int64_t s = cdiv(2 * 148, tiles);
  s = std::min<int64_t>(s, cdiv(B, 2 * DT));
  return std::max<int64_t>(1, std::min<int64_t>(s, 32));
}

template <typename K>
static int set_smem(K kernel, size_t bytes, const char* name) {
  if (bytes > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
    if (e != cudaSuccess) {
      set_error("%s: cannot reserve %zu B of shared memory: %s", name, bytes, cudaGetErrorString(e));
      return SCVI_ECUDA;
    }
  }
  return SCVI_OK;
}

}  // namespace scvi

namespace scvi {
int64_t tc_decoder_fwd_ws_bytes(int64_t B, int64_t H, int64_t G, int likelihood, bool aops);
int tc_decoder_nll_fwd(const float* h, int64_t ldh, int64_t B, int64_t H, int64_t G, const float* Ws, const float* bs,
                       const float* Wp, const float* bp, const float* px_r, const void* x, int64_t ldx,
                       const int64_t* row_idx, const float* library, int likelihood, float* reconst,
                       float* row_stats, const float* g_reconst, float* dh, int64_t lddh, void* ws, cudaStream_t st,
                       bool prepacked, int xkind);
int tc_decoder_nll_prepack(int64_t B, int64_t H, int64_t G, const float* Ws, const float* bs, const float* Wp,
                           const float* bp, const float* px_r, int likelihood, bool want_dh, void* ws, cudaStream_t st);
int64_t tc_decoder_bwd_ws_bytes(int64_t B, int64_t G, int likelihood);
int tc_decoder_nll_bwd_dw(const float* h, int64_t ldh, int64_t B, int64_t G, const float* Ws, const float* bs,
                          const float* Wp, const float* bp, const float* px_r, const void* x, int64_t ldx,
                          const int64_t* row_idx, const float* row_stats, const float* g_reconst, int likelihood,
                          float* dWs, float* dbs, float* dWp, float* dbp, float* dpx_r, void* ws, const void* fwd_ws,
                          int xkind, const float* library, cudaStream_t st);
// count-operand kind carried in the `precision` argument (scvi_b200.h: SCVI_X_INTEGER / SCVI_X_U8)
// bit 4 of the result: SCVI_PREC_AOPS (staged operands kept by the forward / consumed by the backward)
static inline int take_xkind(int& precision) {
  int k = (precision & SCVI_X_U8) ? 2 : (precision & SCVI_X_INTEGER) ? 1 : 0;
  if ((precision & SCVI_PREC_AOPS) && k != 0) k |= 16;
  precision &= ~(SCVI_X_U8 | SCVI_X_INTEGER | SCVI_PREC_AOPS);
  return k;
}
}  // namespace scvi

extern "C" {

int64_t scvi_decoder_nll_fwd_ws_bytes(int64_t B, int64_t H, int64_t G, int likelihood, int precision) {
  const int xk = scvi::take_xkind(precision);
  if (precision == SCVI_PREC_TC) return scvi::tc_decoder_fwd_ws_bytes(B, H, G, likelihood, (xk & 16) != 0);
  return 0;
}

int scvi_decoder_nll_fwd(const float* h, int64_t ldh, int64_t B, int64_t H, int64_t G, const float* Ws,
                         const float* bs, const float* Wp, const float* bp, const float* px_r, const void* xv,
                         int64_t ldx, const int64_t* row_idx, const float* library, int likelihood, int precision,
                         float* reconst, float* row_stats, const float* g_reconst, float* dh, int64_t lddh, void* ws,
                         void* stream) {
  using namespace scvi;
  const float* x = reinterpret_cast<const float*>(xv);
  const int xkind_full = take_xkind(precision);
  const int xkind = xkind_full & 15;
  const bool zinb = likelihood == SCVI_LIK_ZINB;
  SCVI_REQUIRE(likelihood == SCVI_LIK_NB || zinb, "scvi_decoder_nll_fwd: unknown likelihood %d", likelihood);
  SCVI_REQUIRE(h && Ws && bs && px_r && x && library && reconst && row_stats && (!zinb || (Wp && bp)),
               "scvi_decoder_nll_fwd: NULL pointer");
  SCVI_REQUIRE(B >= 0 && G > 0 && H > 0 && H % 4 == 0 && ldh >= H && ldh % 4 == 0 && ldx >= G,
               "scvi_decoder_nll_fwd: bad shape (H %% 4 == 0, ldh %% 4 == 0)");
  SCVI_REQUIRE(((uintptr_t)h % 16 == 0) && ((uintptr_t)Ws % 16 == 0) && (!zinb || (uintptr_t)Wp % 16 == 0) &&
                   ((uintptr_t)row_stats % 16 == 0),
               "scvi_decoder_nll_fwd: h / Ws / Wp / row_stats must be 16-byte aligned");
  if (B == 0) return SCVI_OK;
  const bool prepacked = (precision & SCVI_PREC_PREPACKED) != 0;
  precision &= ~SCVI_PREC_PREPACKED;
  if (precision == SCVI_PREC_TC)
    return tc_decoder_nll_fwd(h, ldh, B, H, G, Ws, bs, Wp, bp, px_r, xv, ldx, row_idx, library, likelihood, reconst,
                              row_stats, g_reconst, dh, lddh, ws, (cudaStream_t)stream, prepacked, xkind_full);
  if (xkind == 2) {
    set_error("scvi_decoder_nll_fwd: SCVI_X_U8 needs SCVI_PREC_TC");
    return SCVI_EUNSUPPORTED;
  }
  if (precision != SCVI_PREC_FP32) {
    set_error("scvi_decoder_nll_fwd: unknown precision %d", precision);
    return SCVI_EUNSUPPORTED;
  }
  const size_t smem = sizeof(float) * ((size_t)DT * (H + 4) * (zinb ? 3 : 2) + DT * 5);
  SCVI_REQUIRE(smem <= 220 * 1024, "scvi_decoder_nll_fwd: n_hidden %lld too large", (long long)H);
  cudaStream_t st = (cudaStream_t)stream;
  const unsigned grid = (unsigned)cdiv(B, DT);
  int rc;
  if (zinb) {
    if ((rc = set_smem(decoder_nll_fwd_kernel<true>, smem, "scvi_decoder_nll_fwd"))) return rc;
    decoder_nll_fwd_kernel<true><<<grid, DTHREADS, smem, st>>>(h, ldh, B, (int)H, G, Ws, bs, Wp, bp, px_r, x, ldx, row_idx, library, 1e-8f, reconst, row_stats);
  } else {
    if ((rc = set_smem(decoder_nll_fwd_kernel<false>, smem, "scvi_decoder_nll_fwd"))) return rc;
    decoder_nll_fwd_kernel<false><<<grid, DTHREADS, smem, st>>>(h, ldh, B, (int)H, G, Ws, bs, Wp, bp, px_r, x, ldx, row_idx, library, 1e-8f, reconst, row_stats);
  }
  SCVI_LAUNCH_CHECK("scvi_decoder_nll_fwd");
  if (dh)  // FFMA path: the dh kernel of the backward, run right after the forward
    return scvi_decoder_nll_bwd(h, ldh, B, H, G, Ws, bs, Wp, bp, px_r, x, ldx, row_idx, library, row_stats, g_reconst,
                                likelihood, precision, dh, lddh, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr,
                                nullptr, stream);
  return SCVI_OK;
}

int scvi_decoder_nll_prepack(int64_t B, int64_t H, int64_t G, const float* Ws, const float* bs, const float* Wp,
                             const float* bp, const float* px_r, int likelihood, int precision, int want_dh,
                             void* ws, void* stream) {
  using namespace scvi;
  const bool zinb = likelihood == SCVI_LIK_ZINB;
  SCVI_REQUIRE(likelihood == SCVI_LIK_NB || zinb, "scvi_decoder_nll_prepack: unknown likelihood %d", likelihood);
  SCVI_REQUIRE(Ws && bs && px_r && (!zinb || (Wp && bp)), "scvi_decoder_nll_prepack: NULL pointer");
  SCVI_REQUIRE(B >= 0 && G > 0 && H > 0, "scvi_decoder_nll_prepack: bad shape");
  if (precision != SCVI_PREC_TC) return SCVI_OK;  // the FFMA path has nothing to prepare
  if (B == 0) return SCVI_OK;
  return tc_decoder_nll_prepack(B, H, G, Ws, bs, Wp, bp, px_r, likelihood, want_dh != 0, ws, (cudaStream_t)stream);
}

int64_t scvi_decoder_nll_bwd_ws_bytes(int64_t B, int64_t H, int64_t G, int likelihood) {
  const int64_t S = scvi::dw_row_splits(B, G);
  const int64_t nheads = likelihood == SCVI_LIK_ZINB ? 2 : 1;
  const int64_t ffma = 4 * (S * nheads * G * H + S * 3 * G) + 256;
  return H == 128 ? std::max<int64_t>(ffma, scvi::tc_decoder_bwd_ws_bytes(B, G, likelihood)) : ffma;
}

#define SCVI_DISPATCH_NH4(KERNEL, ZINB, NH4, ...)                                  \
  do {                                                                             \
    if ((rc = set_smem(KERNEL<ZINB, NH4>, smem, name))) return rc;                 \
    KERNEL<ZINB, NH4><<<grid, DTHREADS, smem, st>>>(__VA_ARGS__);                  \
  } while (0)

#define SCVI_DISPATCH(KERNEL, ...)                                                 \
  do {                                                                             \
    if (zinb) {                                                                    \
      if (H == 64) SCVI_DISPATCH_NH4(KERNEL, true, 1, __VA_ARGS__);                \
      else if (H == 128) SCVI_DISPATCH_NH4(KERNEL, true, 2, __VA_ARGS__);          \
      else SCVI_DISPATCH_NH4(KERNEL, true, 4, __VA_ARGS__);                        \
    } else {                                                                       \
      if (H == 64) SCVI_DISPATCH_NH4(KERNEL, false, 1, __VA_ARGS__);               \
      else if (H == 128) SCVI_DISPATCH_NH4(KERNEL, false, 2, __VA_ARGS__);         \
      else SCVI_DISPATCH_NH4(KERNEL, false, 4, __VA_ARGS__);                       \
    }                                                                              \
  } while (0)

int scvi_decoder_nll_bwd(const float* h, int64_t ldh, int64_t B, int64_t H, int64_t G, const float* Ws,
                         const float* bs, const float* Wp, const float* bp, const float* px_r, const void* xv,
                         int64_t ldx, const int64_t* row_idx, const float* library, const float* row_stats,
                         const float* g_reconst, int likelihood, int precision, float* dh, int64_t lddh,
                         float* dWs, float* dbs, float* dWp, float* dbp, float* dpx_r, void* ws, const void* fwd_ws,
                         void* stream) {
  using namespace scvi;
  const char* name = "scvi_decoder_nll_bwd";
  const float* x = reinterpret_cast<const float*>(xv);
  const int xkind_full = take_xkind(precision);
  const int xkind = xkind_full & 15;
  if (xkind == 2 && (dh || precision != SCVI_PREC_TC || H != 128)) {
    set_error("%s: SCVI_X_U8 is read by the tensor-core weight-gradient kernel only (SCVI_PREC_TC, H == 128, dh == NULL)", name);
    return SCVI_EUNSUPPORTED;
  }
  const bool zinb = likelihood == SCVI_LIK_ZINB;
  SCVI_REQUIRE(likelihood == SCVI_LIK_NB || zinb, "%s: unknown likelihood %d", name, likelihood);
  const bool want_dw = dWs || dWp || dbs || dbp || dpx_r;
  SCVI_REQUIRE(h && Ws && bs && px_r && x && row_stats && (ws || !want_dw) && (!zinb || (Wp && bp)),
               "%s: NULL pointer", name);
  SCVI_REQUIRE(B > 0 && G > 0 && ldh >= H && ldh % 4 == 0 && ldx >= G, "%s: bad shape", name);
  if (!(H == 64 || H == 128 || H == 256)) {
    set_error("%s: n_hidden must be 64, 128 or 256 (got %lld)", name, (long long)H);
    return SCVI_EUNSUPPORTED;
  }
  SCVI_REQUIRE(!dh || (lddh >= H && lddh % 4 == 0 && (uintptr_t)dh % 16 == 0), "%s: dh must be 16-byte aligned", name);
  if (precision != SCVI_PREC_FP32 && precision != SCVI_PREC_TC) {
    set_error("%s: unknown precision %d", name, precision);
    return SCVI_EUNSUPPORTED;
  }
  cudaStream_t st = (cudaStream_t)stream;
  int rc;
  if (dh) {
    const size_t smem = sizeof(float) * ((size_t)DT * (H + 4) * (zinb ? 3 : 2) + (size_t)DT * (DT + 4) * (zinb ? 2 : 1) + DT * 5);
    const unsigned grid = (unsigned)cdiv(B, DT);
    SCVI_DISPATCH(decoder_nll_bwd_dh_kernel, h, ldh, B, G, Ws, bs, Wp, bp, px_r, x, ldx, row_idx, row_stats, g_reconst, 1e-8f, dh, lddh);
    SCVI_LAUNCH_CHECK("scvi_decoder_nll_bwd(dh)");
  }
  if (want_dw && precision == SCVI_PREC_TC && H == 128)
    return tc_decoder_nll_bwd_dw(h, ldh, B, G, Ws, bs, Wp, bp, px_r, xv, ldx, row_idx, row_stats, g_reconst, likelihood,
                                 dWs, dbs, dWp, dbp, dpx_r, ws, fwd_ws, xkind_full, library, st);
  if (want_dw) {
    const int64_t S = dw_row_splits(B, G);
    const int nheads = zinb ? 2 : 1;
    float* slabW = (float*)ws;
    float* slabv = slabW + S * nheads * G * H;
    const int64_t rps = cdiv(cdiv(B, S), DT) * DT;
    const size_t smem = sizeof(float) * ((size_t)DT * (H + 4) * (zinb ? 3 : 2) + (size_t)DT * (DT + 4) * (zinb ? 2 : 1) + DT * 5);
    SCVI_REQUIRE((size_t)DT * (DT + 4) >= 3 * 16 * DT, "%s: internal smem layout", name);
    dim3 grid((unsigned)cdiv(G, DT), (unsigned)S);
    SCVI_DISPATCH(decoder_nll_bwd_dw_kernel, h, ldh, B, G, Ws, bs, Wp, bp, px_r, x, ldx, row_idx, row_stats, g_reconst, 1e-8f, rps, slabW, slabv);
    SCVI_LAUNCH_CHECK("scvi_decoder_nll_bwd(dW)");
    decoder_bwd_reduce_kernel<<<(unsigned)cdiv(G * H, 256), 256, 0, st>>>(slabW, slabv, S, nheads, G * H, G, dWs, dWp, dbs, dbp, dpx_r);
    SCVI_LAUNCH_CHECK("scvi_decoder_nll_bwd(reduce)");
  }
  return SCVI_OK;
}

int scvi_decoder_params_fwd(const float* h, int64_t ldh, int64_t B, int64_t H, int64_t G, const float* Ws,
                            const float* bs, const float* Wp, const float* bp, const float* library,
                            float* px_scale, float* px_rate, float* px_dropout, int64_t ldo, void* stream) {
  using namespace scvi;
  SCVI_REQUIRE(h && Ws && bs && (!px_dropout || (Wp && bp)) && (!px_rate || library),
               "scvi_decoder_params_fwd: NULL pointer");
  SCVI_REQUIRE(B >= 0 && G > 0 && H > 0 && H % 4 == 0 && ldh >= H && ldh % 4 == 0 && ldo >= G,
               "scvi_decoder_params_fwd: bad shape");
  if (B == 0) return SCVI_OK;
  const size_t smem = sizeof(float) * ((size_t)DT * (H + 4) * 2 + DT);
  int rc;
  if ((rc = set_smem(decoder_params_kernel, smem, "scvi_decoder_params_fwd"))) return rc;
  decoder_params_kernel<<<(unsigned)cdiv(B, DT), DTHREADS, smem, (cudaStream_t)stream>>>(h, ldh, B, (int)H, G, Ws, bs, Wp, bp, library, px_scale, px_rate, px_dropout, ldo);
  SCVI_LAUNCH_CHECK("scvi_decoder_params_fwd");
  return SCVI_OK;
}
}
