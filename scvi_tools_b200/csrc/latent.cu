// Latent heads of Encoder.forward + reparameterisation + Normal KL, forward and backward.
// Replaces (reference src/scvi/nn/_base_components.py:311-317)
//   q_m = mean_encoder(q); q_v = exp(var_encoder(q)) + var_eps; z = Normal(q_m, sqrt(q_v)).rsample()
// and kl_divergence(qz, Normal(0,1)).sum(-1) (src/scvi/module/_vae.py:563-565), where torch's
// kl_normal_normal gives 0.5 * (s^2 + q_m^2 - 1 - log(s^2)) with s = sqrt(q_v).
#include "common.cuh"

namespace scvi {

constexpr int LAT_WARPS = 4;
constexpr int LAT_RPW_MAX = 4;  // rows a warp carries through the head GEMV together when the batch is large

// one warp per LAT_RPW rows; weights staged in smem as Wt[k][2L] (k-major => lane j reads Wt[k][j]), the rows'
// activations as arT[k][LAT_RPW] so that one broadcast float4 feeds four independent FMA chains per lane.
// LAT_RPW = 4 for tall inputs (SCANVI's label-enumerated 20 x B rows); 1 for few-cell minibatches, where the
// number of warps in flight matters more than the instruction count (measured: B = 1024, 64 CTAs of 4 rows per
// warp 18 us, 256 CTAs of 1 row per warp 11.5 us).
template <int LAT_RPW>
__global__ void __launch_bounds__(LAT_WARPS * 32)
    latent_fwd_kernel(const float* __restrict__ a, int64_t lda, int64_t B, int H, int L,
                      const float* __restrict__ Wm, const float* __restrict__ bm, const float* __restrict__ Wv,
                      const float* __restrict__ bv, const float* __restrict__ noise, float var_eps,
                      float* __restrict__ q_m, float* __restrict__ q_v, float* __restrict__ z,
                      float* __restrict__ kl) {
  extern __shared__ float sm[];
  const int L2 = 2 * L;
  float* Wt = sm;                                                  // [H][L2]
  static_assert(LAT_RPW == 1 || LAT_RPW == 4, "rows per warp: 1 or 4");
  float* arT = sm + (((size_t)H * L2 + 3) & ~(size_t)3);           // [LAT_WARPS][H][LAT_RPW], 16-byte aligned
  float* res = arT + (size_t)LAT_WARPS * H * LAT_RPW;              // [LAT_WARPS][LAT_RPW][L2]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int j = 0; j < L2; ++j) {
    const float* src = (j < L) ? Wm + (size_t)j * H : Wv + (size_t)(j - L) * H;
    for (int k = tid; k < H; k += blockDim.x) Wt[k * L2 + j] = src[k];
  }
  __syncthreads();
  float* art = arT + (size_t)warp * H * LAT_RPW;
  float* rs = res + (size_t)warp * LAT_RPW * L2;
  for (int64_t rb = ((int64_t)blockIdx.x * LAT_WARPS + warp) * LAT_RPW; rb < B;
       rb += (int64_t)gridDim.x * LAT_WARPS * LAT_RPW) {
#pragma unroll
    for (int rr = 0; rr < LAT_RPW; ++rr)
      for (int k = lane; k < H; k += 32) art[k * LAT_RPW + rr] = (rb + rr < B) ? a[(rb + rr) * lda + k] : 0.f;
    __syncwarp();
    for (int j = lane; j < L2; j += 32) {
      const float b0 = (j < L) ? bm[j] : bv[j - L];
      float s0 = b0, s1 = b0, s2 = b0, s3 = b0;
      for (int k = 0; k < H; ++k) {
        const float w = Wt[k * L2 + j];
        if (LAT_RPW == 4) {
          const float4 av = *reinterpret_cast<const float4*>(art + k * LAT_RPW);
          s0 = fmaf(av.x, w, s0);
          s1 = fmaf(av.y, w, s1);
          s2 = fmaf(av.z, w, s2);
          s3 = fmaf(av.w, w, s3);
        } else {
          s0 = fmaf(art[k], w, s0);
        }
      }
      rs[0 * L2 + j] = s0;
      if (LAT_RPW == 4) {
        rs[1 * L2 + j] = s1;
        rs[2 * L2 + j] = s2;
        rs[3 * L2 + j] = s3;
      }
    }
    __syncwarp();
    for (int rr = 0; rr < LAT_RPW; ++rr) {
      const int64_t r = rb + rr;
      if (r >= B) break;  // warp-uniform
      float klp = 0.f;
      for (int l = lane; l < L; l += 32) {
        const float m = rs[rr * L2 + l];
        const float v = expf(rs[rr * L2 + L + l]) + var_eps;
        const float s = sqrtf(v);
        const float s2 = s * s;
        q_m[r * L + l] = m;
        q_v[r * L + l] = v;
        z[r * L + l] = fmaf(s, noise[r * L + l], m);
        klp += 0.5f * (s2 + m * m - 1.f - logf(s2));
      }
      klp = warp_sum(klp);
      if (lane == 0) kl[r] = klp;
    }
    __syncwarp();
  }
}

constexpr int LATB_ROWS = 16;
constexpr int LATB_THREADS = 256;  // launch bound; 128 threads when H <= 128 (one hidden column per thread)

__global__ void __launch_bounds__(LATB_THREADS)
    latent_bwd_kernel(const float* __restrict__ a, int64_t lda, int64_t B, int H, int L,
                      const float* __restrict__ Wm, const float* __restrict__ Wv, const float* __restrict__ noise,
                      float var_eps, const float* __restrict__ q_m, const float* __restrict__ q_v,
                      const float* __restrict__ g_qm, const float* __restrict__ g_qv, const float* __restrict__ g_z,
                      const float* __restrict__ g_kl, float* __restrict__ da, int64_t ldda,
                      float* __restrict__ partW /*[ctas][2L][H]*/, float* __restrict__ partb /*[ctas][2L]*/) {
  extern __shared__ float sm[];
  const int L2 = 2 * L;
  float* Wn = sm;                            // [L2][H]   rows 0..L-1 = Wm, L..2L-1 = Wv
  float* as = Wn + (size_t)L2 * H;           // [LATB_ROWS][H]
  float* dqT = sm + (((size_t)L2 * H + (size_t)LATB_ROWS * H + 3) & ~(size_t)3);  // [L2][LATB_ROWS]  (dq_m | dt), 16-byte aligned
  const int tid = threadIdx.x;
  const int64_t r0 = (int64_t)blockIdx.x * LATB_ROWS;
  const int cnt = (int)min((int64_t)LATB_ROWS, B - r0);
  for (int e = tid; e < L2 * H; e += blockDim.x) Wn[e] = (e < L * H) ? Wm[e] : Wv[e - L * H];
  for (int r = 0; r < LATB_ROWS; ++r)
    for (int k = tid; k < H; k += blockDim.x) as[r * H + k] = (r < cnt) ? a[(r0 + r) * lda + k] : 0.f;
  for (int e = tid; e < LATB_ROWS * L; e += blockDim.x) {
    const int r = e / L, l = e % L;
    float dm = 0.f, dt = 0.f;
    if (r < cnt) {
      const int64_t i = (r0 + r) * L + l;
      const float m = q_m[i], v = q_v[i];
      const float gk = g_kl ? g_kl[r0 + r] : 0.f;
      const float gz = g_z ? g_z[i] : 0.f;
      dm = (g_qm ? g_qm[i] : 0.f) + gz + gk * m;
      const float s = sqrtf(v);
      float dv = (g_qv ? g_qv[i] : 0.f) + gz * noise[i] / (2.f * s) + gk * 0.5f * (1.f - 1.f / v);
      dt = dv * (v - var_eps);  // d exp(t) = exp(t) = q_v - var_eps
    }
    dqT[l * LATB_ROWS + r] = dm;  // rows past the batch hold zeros: the loops below run over all LATB_ROWS
    dqT[(L + l) * LATB_ROWS + r] = dt;
  }
  __syncthreads();
  static_assert(LATB_ROWS == 16, "the register blocking below is written for 16 rows");
  for (int k = tid; k < H; k += blockDim.x) {
    // da[r][k] = sum_j dq[r][j] * Wn[j][k]: one W value and four broadcast float4 of dq^T per j, 16 running sums
    float s[LATB_ROWS];
#pragma unroll
    for (int r = 0; r < LATB_ROWS; ++r) s[r] = 0.f;
    for (int j = 0; j < L2; ++j) {
      const float w = Wn[j * H + k];
      const float4* q4 = reinterpret_cast<const float4*>(dqT + j * LATB_ROWS);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float4 d = q4[q];
        s[4 * q + 0] = fmaf(d.x, w, s[4 * q + 0]);
        s[4 * q + 1] = fmaf(d.y, w, s[4 * q + 1]);
        s[4 * q + 2] = fmaf(d.z, w, s[4 * q + 2]);
        s[4 * q + 3] = fmaf(d.w, w, s[4 * q + 3]);
      }
    }
#pragma unroll
    for (int r = 0; r < LATB_ROWS; ++r)
      if (r < cnt) da[(r0 + r) * ldda + k] = s[r];
    // dW[j][k] partial = sum_r dq[r][j] * a[r][k]: the 16 activations of column k stay in registers
    float av[LATB_ROWS];
#pragma unroll
    for (int r = 0; r < LATB_ROWS; ++r) av[r] = as[r * H + k];
#pragma unroll 2
    for (int j = 0; j < L2; ++j) {
      const float4* q4 = reinterpret_cast<const float4*>(dqT + j * LATB_ROWS);
      float t = 0.f;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float4 d = q4[q];
        t = fmaf(d.x, av[4 * q + 0], t);
        t = fmaf(d.y, av[4 * q + 1], t);
        t = fmaf(d.z, av[4 * q + 2], t);
        t = fmaf(d.w, av[4 * q + 3], t);
      }
      partW[((int64_t)blockIdx.x * L2 + j) * H + k] = t;
    }
  }
  for (int j = tid; j < L2; j += blockDim.x) {
    float t = 0.f;
    for (int r = 0; r < cnt; ++r) t += dqT[j * LATB_ROWS + r];
    partb[(int64_t)blockIdx.x * L2 + j] = t;
  }
}

// ---- small per-row reductions of the semi-supervised / multi-modal losses ---------------------------------
// out[r] = sum_d log N(value[r,d]; loc[r,d], scale[r,d])   (torch.distributions.Normal.log_prob(value).sum(-1);
// `second` is the variance instead of the scale when is_var).  One thread per row, fixed summation order.
__global__ void normal_logprob_sum_fwd_kernel(const float* __restrict__ value, const float* __restrict__ loc,
                                              const float* __restrict__ second, int64_t R, int D, int is_var,
                                              float* __restrict__ out) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= R) return;
  const float c = 0.9189385332046727f;  // log(sqrt(2 pi))
  float acc = 0.f;
  for (int d = 0; d < D; ++d) {
    const float v = value[r * D + d], m = loc[r * D + d], s2 = second[r * D + d];
    const float var = is_var ? s2 : s2 * s2;
    const float log_scale = is_var ? 0.5f * logf(s2) : logf(s2);
    const float df = v - m;
    acc += -(df * df) / (2.f * var) - log_scale - c;
  }
  out[r] = acc;
}

// gradients of the above times gout[r]; any output may be NULL
__global__ void normal_logprob_sum_bwd_kernel(const float* __restrict__ value, const float* __restrict__ loc,
                                              const float* __restrict__ second, const float* __restrict__ gout,
                                              int64_t R, int D, int is_var, float* __restrict__ dvalue,
                                              float* __restrict__ dloc, float* __restrict__ dsecond) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= R * D) return;
  const float g = gout[e / D];
  const float v = value[e], m = loc[e], s2 = second[e];
  const float df = v - m;
  const float var = is_var ? s2 : s2 * s2;
  const float dv = -g * df / var;
  if (dvalue) dvalue[e] = dv;
  if (dloc) dloc[e] = -dv;
  if (dsecond) {
    if (is_var)
      dsecond[e] = g * (df * df / (2.f * var * var) - 0.5f / var);
    else
      dsecond[e] = g * (df * df / (var * s2) - 1.f / s2);
  }
}

// KL(Categorical(probs = p) || Categorical(probs = q)) per row, torch's construction: both arguments are
// normalised by their row sum, logits = log(clamp(probs, eps, 1 - eps)), t = p (log p - log q), t = inf where
// q == 0, then t = 0 where p == 0, summed over the classes.  q: [R, C] (ldq = C) or one row for all (ldq = 0).
// With dp != NULL the same pass writes gout[r] * d(kl[r]) / d(p[r, :]) (q carries no gradient).
__global__ void categorical_kl_kernel(const float* __restrict__ p, const float* __restrict__ q, int64_t ldq,
                                      int64_t R, int C, const float* __restrict__ gout, float* __restrict__ kl,
                                      float* __restrict__ dp) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= R) return;
  const float eps = 1.1920928955078125e-07f;  // torch.finfo(float32).eps
  const float* pr = p + r * C;
  const float* qr = q + r * ldq;
  float sp = 0.f, sq = 0.f;
  for (int c = 0; c < C; ++c) {
    sp += pr[c];
    sq += qr[c];
  }
  float acc = 0.f, dot = 0.f;
  for (int c = 0; c < C; ++c) {
    const float pn = pr[c] / sp, qn = qr[c] / sq;
    const float lp = logf(fminf(fmaxf(pn, eps), 1.f - eps));
    const float lq = logf(fminf(fmaxf(qn, eps), 1.f - eps));
    float t = pn * (lp - lq);
    if (qn == 0.f) t = INFINITY;
    if (pn == 0.f) t = 0.f;
    acc += t;
    if (dp) {
      float dpn = 0.f;
      if (qn != 0.f && pn != 0.f) dpn = (lp - lq) + ((pn >= eps && pn <= 1.f - eps) ? 1.f : 0.f);
      dot += dpn * pn;
    }
  }
  if (kl) kl[r] = acc;
  if (dp) {
    const float g = gout[r] / sp;
    for (int c = 0; c < C; ++c) {
      const float pn = pr[c] / sp, qn = qr[c] / sq;
      const float lp = logf(fminf(fmaxf(pn, eps), 1.f - eps));
      const float lq = logf(fminf(fmaxf(qn, eps), 1.f - eps));
      float dpn = 0.f;
      if (qn != 0.f && pn != 0.f) dpn = (lp - lq) + ((pn >= eps && pn <= 1.f - eps) ? 1.f : 0.f);
      dp[r * C + c] = g * (dpn - dot);
    }
  }
}

}  // namespace scvi

extern "C" {

int scvi_latent_fwd(const float* a, int64_t lda, int64_t B, int64_t H, int64_t L, const float* Wm, const float* bm,
                    const float* Wv, const float* bv, const float* noise, float var_eps, float* q_m, float* q_v,
                    float* z, float* kl, void* stream) {
  using namespace scvi;
  SCVI_REQUIRE(a && Wm && bm && Wv && bv && noise && q_m && q_v && z && kl, "scvi_latent_fwd: NULL pointer");
  SCVI_REQUIRE(B >= 0 && H > 0 && L > 0 && L <= 32 && lda >= H, "scvi_latent_fwd: bad shape (L <= 32)");
  if (B == 0) return SCVI_OK;
  const int rpw = (cdiv(B, LAT_WARPS * LAT_RPW_MAX) >= 2 * 148) ? LAT_RPW_MAX : 1;
  const size_t smem = sizeof(float) * ((size_t)H * 2 * L + 4 + (size_t)LAT_WARPS * rpw * (H + 2 * L));
  static std::atomic<uint64_t> optin1{0}, optin4{0};
  if (smem > 48 * 1024) {
    SCVI_REQUIRE(smem <= 200 * 1024, "scvi_latent_fwd: H*L too large for shared memory");
    if (int rc = rpw == 1 ? smem_optin((const void*)latent_fwd_kernel<1>, 200 * 1024, optin1, "scvi_latent_fwd")
                          : smem_optin((const void*)latent_fwd_kernel<LAT_RPW_MAX>, 200 * 1024, optin4, "scvi_latent_fwd"))
      return rc;
  }
  const unsigned blocks = (unsigned)std::min<int64_t>(cdiv(B, LAT_WARPS * rpw), 148 * 4);
  if (rpw == 1)
    latent_fwd_kernel<1><<<blocks, LAT_WARPS * 32, smem, (cudaStream_t)stream>>>(a, lda, B, (int)H, (int)L, Wm, bm, Wv, bv, noise, var_eps, q_m, q_v, z, kl);
  else
    latent_fwd_kernel<LAT_RPW_MAX><<<blocks, LAT_WARPS * 32, smem, (cudaStream_t)stream>>>(a, lda, B, (int)H, (int)L, Wm, bm, Wv, bv, noise, var_eps, q_m, q_v, z, kl);
  SCVI_LAUNCH_CHECK("scvi_latent_fwd");
  return SCVI_OK;
}

int64_t scvi_latent_bwd_ws_bytes(int64_t B, int64_t H, int64_t L) {
  const int64_t ctas = scvi::cdiv(B, scvi::LATB_ROWS);
  return 4 * (ctas * 2 * L * H + ctas * 2 * L) + 256;
}

int scvi_latent_bwd(const float* a, int64_t lda, int64_t B, int64_t H, int64_t L, const float* Wm, const float* Wv,
                    const float* noise, float var_eps, const float* q_m, const float* q_v, const float* g_qm,
                    const float* g_qv, const float* g_z, const float* g_kl, float* da, int64_t ldda, float* dWm,
                    float* dbm, float* dWv, float* dbv, void* ws, void* stream) {
  using namespace scvi;
  SCVI_REQUIRE(a && Wm && Wv && noise && q_m && q_v && da && ws, "scvi_latent_bwd: NULL pointer");
  SCVI_REQUIRE(B > 0 && H > 0 && L > 0 && L <= 32 && lda >= H && ldda >= H, "scvi_latent_bwd: bad shape");
  const size_t smem = sizeof(float) * ((size_t)2 * L * H + (size_t)LATB_ROWS * H + (size_t)LATB_ROWS * 2 * L + 4);
  static std::atomic<uint64_t> optin{0};
  if (smem > 48 * 1024) {
    SCVI_REQUIRE(smem <= 200 * 1024, "scvi_latent_bwd: H*L too large for shared memory");
    if (int rc = smem_optin((const void*)latent_bwd_kernel, 200 * 1024, optin, "scvi_latent_bwd")) return rc;
  }
  const int64_t ctas = cdiv(B, LATB_ROWS);
  float* partW = (float*)ws;
  float* partb = partW + ctas * 2 * L * H;
  cudaStream_t st = (cudaStream_t)stream;
  latent_bwd_kernel<<<(unsigned)ctas, H > 128 ? LATB_THREADS : 128, smem, st>>>(a, lda, B, (int)H, (int)L, Wm, Wv, noise, var_eps, q_m, q_v, g_qm, g_qv, g_z, g_kl, da, ldda, partW, partb);
  SCVI_LAUNCH_CHECK("scvi_latent_bwd");
  {
    ColsumSegs segs{};
    segs.n = 2;
    segs.dst[0] = dWm; segs.begin[0] = 0; segs.end[0] = L * H;
    segs.dst[1] = dWv; segs.begin[1] = L * H; segs.end[1] = 2 * L * H;
    int rc = launch_colsum(partW, ctas, 2 * L * H, 2 * L * H, segs, st, "scvi_latent_bwd(reduce W)");
    if (rc) return rc;
    segs.dst[0] = dbm; segs.end[0] = L;
    segs.dst[1] = dbv; segs.begin[1] = L; segs.end[1] = 2 * L;
    rc = launch_colsum(partb, ctas, 2 * L, 2 * L, segs, st, "scvi_latent_bwd(reduce b)");
    if (rc) return rc;
  }
  return SCVI_OK;
}

int scvi_normal_logprob_sum_fwd(const float* value, const float* loc, const float* second, int64_t R, int64_t D,
                                int is_var, float* out, void* stream) {
  using namespace scvi;
  SCVI_REQUIRE(R >= 0 && D > 0 && D <= 4096, "scvi_normal_logprob_sum_fwd: bad shape");
  if (R == 0) return SCVI_OK;
  SCVI_REQUIRE(value && loc && second && out, "scvi_normal_logprob_sum_fwd: NULL pointer");
  normal_logprob_sum_fwd_kernel<<<(unsigned)cdiv(R, 128), 128, 0, (cudaStream_t)stream>>>(value, loc, second, R, (int)D, is_var, out);
  SCVI_LAUNCH_CHECK("scvi_normal_logprob_sum_fwd");
  return SCVI_OK;
}

int scvi_normal_logprob_sum_bwd(const float* value, const float* loc, const float* second, const float* gout,
                                int64_t R, int64_t D, int is_var, float* dvalue, float* dloc, float* dsecond,
                                void* stream) {
  using namespace scvi;
  SCVI_REQUIRE(R >= 0 && D > 0 && D <= 4096, "scvi_normal_logprob_sum_bwd: bad shape");
  if (R == 0) return SCVI_OK;
  SCVI_REQUIRE(value && loc && second && gout, "scvi_normal_logprob_sum_bwd: NULL pointer");
  normal_logprob_sum_bwd_kernel<<<(unsigned)cdiv(R * D, 256), 256, 0, (cudaStream_t)stream>>>(value, loc, second, gout, R, (int)D, is_var, dvalue, dloc, dsecond);
  SCVI_LAUNCH_CHECK("scvi_normal_logprob_sum_bwd");
  return SCVI_OK;
}

int scvi_categorical_kl(const float* p, const float* q, int64_t ldq, int64_t R, int64_t C, const float* gout,
                        float* kl, float* dp, void* stream) {
  using namespace scvi;
  SCVI_REQUIRE(R >= 0 && C > 0 && C <= 65536 && (ldq == 0 || ldq >= C), "scvi_categorical_kl: bad shape");
  if (R == 0) return SCVI_OK;
  SCVI_REQUIRE(p && q && (kl || dp) && (!dp || gout), "scvi_categorical_kl: NULL pointer");
  categorical_kl_kernel<<<(unsigned)cdiv(R, 128), 128, 0, (cudaStream_t)stream>>>(p, q, ldq, R, (int)C, gout, kl, dp);
  SCVI_LAUNCH_CHECK("scvi_categorical_kl");
  return SCVI_OK;
}
}
