// Standalone NB / ZINB log-likelihood kernels (forward + backward), elementwise over [B,G].
// Replace log_nb_positive / log_zinb_positive as called by NegativeBinomial.log_prob /
// ZeroInflatedNegativeBinomial.log_prob (reference
// src/scvi/distributions/_negative_binomial.py:36-135,445-459,573-586).
// HBM-bound.  Algorithmic bytes: NB fwd 12*B*G + 4*G, ZINB fwd 16*B*G + 4*G,
// NB bwd 20*B*G, ZINB bwd 28*B*G (+ dtheta partials) -- SURVEY.md section 8(d).
#include "fastmath.cuh"
#include "likelihood.cuh"

namespace scvi {

constexpr int LP_THREADS = 256;

// grid: x over gene blocks (LP_THREADS genes, one per thread => per-gene constants hoisted to
// registers), y over row tiles of SCVI_BN_TILE rows.  Coalesced: a warp reads 128 B per row.
template <bool ZINB, bool THETA_PER_GENE>
__global__ void __launch_bounds__(LP_THREADS)
    logprob_fwd_kernel(const float* __restrict__ x, int64_t ldx, const float* __restrict__ mu, int64_t ldmu,
                       const float* __restrict__ theta, int64_t theta_ld, const float* __restrict__ pi,
                       int64_t ldpi, int64_t B, int64_t G, float eps, float* __restrict__ out, int64_t ldo,
                       int32_t* __restrict__ bad_support) {
  const int64_t g = (int64_t)blockIdx.x * LP_THREADS + threadIdx.x;
  const int64_t r0 = (int64_t)blockIdx.y * SCVI_BN_TILE;
  const int64_t r1 = min(r0 + (int64_t)SCVI_BN_TILE, B);
  if (g >= G) return;
  GeneConst c;
  if (THETA_PER_GENE) c = make_gene_const(theta[g], eps);
  bool bad = false;
#pragma unroll 4
  for (int64_t r = r0; r < r1; ++r) {
    const float xv = ld_stream(x + r * ldx + g);
    const float m = ld_stream(mu + r * ldmu + g);
    if (!THETA_PER_GENE) c = make_gene_const(theta[r * theta_ld + g], eps);
    float ll;
    if (ZINB)
      ll = zinb_ll(xv, m, ld_stream(pi + r * ldpi + g), c, eps);
    else
      ll = nb_ll(xv, m, c, eps);
    bad |= (xv < 0.f) || (xv != floorf(xv));
    out[r * ldo + g] = ll;
  }
  if (bad_support != nullptr && bad) *bad_support = 1;
}

template <bool ZINB, bool THETA_PER_GENE>
__global__ void __launch_bounds__(LP_THREADS)
    logprob_bwd_kernel(const float* __restrict__ x, int64_t ldx, const float* __restrict__ mu, int64_t ldmu,
                       const float* __restrict__ theta, int64_t theta_ld, const float* __restrict__ pi,
                       int64_t ldpi, const float* __restrict__ gout, int64_t ldg, int64_t B, int64_t G,
                       float eps, float* __restrict__ dmu, float* __restrict__ dtheta, float* __restrict__ dpi) {
  const int64_t g = (int64_t)blockIdx.x * LP_THREADS + threadIdx.x;
  const int64_t r0 = (int64_t)blockIdx.y * SCVI_BN_TILE;
  const int64_t r1 = min(r0 + (int64_t)SCVI_BN_TILE, B);
  if (g >= G) return;
  GeneConst c;
  if (THETA_PER_GENE) c = make_gene_const(theta[g], eps);
  float dth_acc = 0.f;
#pragma unroll 2
  for (int64_t r = r0; r < r1; ++r) {
    const float xv = ld_stream(x + r * ldx + g);
    const float m = ld_stream(mu + r * ldmu + g);
    const float go = ld_stream(gout + r * ldg + g);
    if (!THETA_PER_GENE) c = make_gene_const(theta[r * theta_ld + g], eps);
    float a, b, d = 0.f;
    if (ZINB)
      zinb_ll_grad(xv, m, ld_stream(pi + r * ldpi + g), c, eps, a, b, d);
    else
      nb_ll_grad(xv, m, c, eps, a, b);
    if (dmu) dmu[r * G + g] = go * a;
    if (ZINB && dpi) dpi[r * G + g] = go * d;
    if (dtheta) {
      if (THETA_PER_GENE)
        dth_acc += go * b;
      else
        dtheta[r * G + g] = go * b;
    }
  }
  if (THETA_PER_GENE && dtheta) dtheta[(int64_t)blockIdx.y * G + g] = dth_acc;  // per-row-tile partial
}

// ---- vectorised fast path: per-gene theta, 16-byte aligned rows -------------------------------------------
// thread = 4 consecutive genes (one 16-byte column) x a tile of SCVI_BN_TILE rows; per-gene constants of the 4
// genes live in registers; hand-written SFU math (fastmath.cuh) keeps the kernels on the HBM roofline:
// NB forward 6 SFU ops per element, ZINB 9.
constexpr int LP4_THREADS = 128;
// rows per straight-line block of the forward kernel (two blocks live in registers): NB reads two operands per
// element, ZINB three -- the same bytes in flight per thread either way
template <bool ZINB>
struct Lp4Rb {
  static constexpr int value = ZINB ? 3 : 4;
};

__device__ __forceinline__ fm::GeneK lp_gene_k(float theta, float eps) {
  fm::GeneK k;
  k.theta = theta;
  k.log_theta = logf(theta + eps);
  k.lg_theta = lgammaf(theta);
  k.rth = 1.f / (theta + eps);
  k.dg_theta = digamma_pos(theta);
  return k;
}

template <bool ZINB>
__global__ void __launch_bounds__(LP4_THREADS, 4)
    logprob_fwd4_kernel(const float* __restrict__ x, int64_t ldx, const float* __restrict__ mu, int64_t ldmu,
                        const float* __restrict__ theta, const float* __restrict__ pi, int64_t ldpi, int64_t B,
                        int64_t G, float eps, float* __restrict__ out, int64_t ldo, int32_t* __restrict__ bad_support,
                        int row_tile) {
  constexpr int LP4_RB = Lp4Rb<ZINB>::value;
  __shared__ float s_lgx1[fm::LGX1_TABLE];
  for (int i = threadIdx.x; i < fm::LGX1_TABLE; i += LP4_THREADS) s_lgx1[i] = lgammaf((float)i + 1.f);
  __syncthreads();
  const int64_t g = ((int64_t)blockIdx.x * LP4_THREADS + threadIdx.x) * 4;
  const int64_t r0 = (int64_t)blockIdx.y * row_tile;
  const int64_t r1 = min(r0 + (int64_t)row_tile, B);
  if (g >= G) return;  // G % 4 == 0 on this path
  const float4 th = *reinterpret_cast<const float4*>(theta + g);
  const fm::GeneK k0 = lp_gene_k(th.x, eps), k1 = lp_gene_k(th.y, eps), k2 = lp_gene_k(th.z, eps), k3 = lp_gene_k(th.w, eps);
  bool bad = false;
  // Blocks of LP4_RB rows, register double-buffered: the loads of block i + 1 are issued BEFORE block i is evaluated,
  // so a warp always has a block in flight (the kernel was waiting on memory: long-scoreboard stalls 5.3 per issue).
  // Per block ONE test that every count indexes the lgamma(x + 1) table, then LP4_RB x 4 straight-line elements
  // (their SFU chains interleave); a block holding anything else -- a count >= 256, a non-integer, a negative --
  // hands the rest of the tile to the generic per-element loop below, which also raises the support flag.
  auto load_block = [&](float4 (&xv)[LP4_RB], float4 (&m)[LP4_RB], float4 (&p)[LP4_RB], int64_t r) {
#pragma unroll
    for (int i = 0; i < LP4_RB; ++i) {
      xv[i] = ld_stream4(x + (r + i) * ldx + g);
      m[i] = ld_stream4(mu + (r + i) * ldmu + g);
      p[i] = ZINB ? ld_stream4(pi + (r + i) * ldpi + g) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
  };
  auto eval_block = [&](const float4 (&xv)[LP4_RB], const float4 (&m)[LP4_RB], const float4 (&p)[LP4_RB], int64_t r) {
    bool ok = true;
#pragma unroll
    for (int i = 0; i < LP4_RB; ++i) ok &= fm::counts_in_table4(xv[i].x, xv[i].y, xv[i].z, xv[i].w);
    if (!ok) return false;
#pragma unroll
    for (int i = 0; i < LP4_RB; ++i) {
      float4 o;
      o.x = fm::ll_mu_count<ZINB>(xv[i].x, m[i].x, p[i].x, k0, eps, s_lgx1);
      o.y = fm::ll_mu_count<ZINB>(xv[i].y, m[i].y, p[i].y, k1, eps, s_lgx1);
      o.z = fm::ll_mu_count<ZINB>(xv[i].z, m[i].z, p[i].z, k2, eps, s_lgx1);
      o.w = fm::ll_mu_count<ZINB>(xv[i].w, m[i].w, p[i].w, k3, eps, s_lgx1);
      *reinterpret_cast<float4*>(out + (r + i) * ldo + g) = o;
    }
    return true;
  };
  int64_t r = r0;
  if (r + LP4_RB <= r1) {
    float4 xA[LP4_RB], mA[LP4_RB], pA[LP4_RB], xB[LP4_RB], mB[LP4_RB], pB[LP4_RB];
    load_block(xA, mA, pA, r);
    while (true) {
      const bool moreB = r + 2 * LP4_RB <= r1;
      if (moreB) load_block(xB, mB, pB, r + LP4_RB);
      if (!eval_block(xA, mA, pA, r)) break;
      r += LP4_RB;
      if (!moreB) break;
      const bool moreA = r + 2 * LP4_RB <= r1;
      if (moreA) load_block(xA, mA, pA, r + LP4_RB);
      if (!eval_block(xB, mB, pB, r)) break;
      r += LP4_RB;
      if (!moreA) break;
    }
  }
  for (; r < r1; ++r) {  // generic per-element path: the rest of a tile after a block that failed the test, and the
                         // rows left over when the tile is not a multiple of the block
    const float4 xv = ld_stream4(x + r * ldx + g);
    const float4 m = ld_stream4(mu + r * ldmu + g);
    float4 p = make_float4(0.f, 0.f, 0.f, 0.f);
    if (ZINB) p = ld_stream4(pi + r * ldpi + g);
    float4 o;
    o.x = fm::ll_mu<ZINB>(xv.x, m.x, p.x, k0, eps, s_lgx1);
    o.y = fm::ll_mu<ZINB>(xv.y, m.y, p.y, k1, eps, s_lgx1);
    o.z = fm::ll_mu<ZINB>(xv.z, m.z, p.z, k2, eps, s_lgx1);
    o.w = fm::ll_mu<ZINB>(xv.w, m.w, p.w, k3, eps, s_lgx1);
    bad |= fm::bad_count(xv.x) | fm::bad_count(xv.y) | fm::bad_count(xv.z) | fm::bad_count(xv.w);
    *reinterpret_cast<float4*>(out + r * ldo + g) = o;
  }
  if (bad_support != nullptr && bad) *bad_support = 1;
}

template <bool ZINB>
__global__ void __launch_bounds__(LP4_THREADS)
    logprob_bwd4_kernel(const float* __restrict__ x, int64_t ldx, const float* __restrict__ mu, int64_t ldmu,
                        const float* __restrict__ theta, const float* __restrict__ pi, int64_t ldpi,
                        const float* __restrict__ gout, int64_t ldg, int64_t B, int64_t G, float eps,
                        float* __restrict__ dmu, float* __restrict__ dtheta, float* __restrict__ dpi) {
  const int64_t g = ((int64_t)blockIdx.x * LP4_THREADS + threadIdx.x) * 4;
  const int64_t r0 = (int64_t)blockIdx.y * SCVI_BN_TILE;
  const int64_t r1 = min(r0 + (int64_t)SCVI_BN_TILE, B);
  if (g >= G) return;
  const float4 th = *reinterpret_cast<const float4*>(theta + g);
  const fm::GeneK k0 = lp_gene_k(th.x, eps), k1 = lp_gene_k(th.y, eps), k2 = lp_gene_k(th.z, eps), k3 = lp_gene_k(th.w, eps);
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 2
  for (int64_t r = r0; r < r1; ++r) {
    const float4 xv = ld_stream4(x + r * ldx + g);
    const float4 m = ld_stream4(mu + r * ldmu + g);
    const float4 go = ld_stream4(gout + r * ldg + g);
    float4 p = make_float4(0.f, 0.f, 0.f, 0.f);
    if (ZINB) p = ld_stream4(pi + r * ldpi + g);
    float4 a, b, d;
    fm::grad_mu<ZINB>(xv.x, m.x, p.x, k0, eps, a.x, b.x, d.x);
    fm::grad_mu<ZINB>(xv.y, m.y, p.y, k1, eps, a.y, b.y, d.y);
    fm::grad_mu<ZINB>(xv.z, m.z, p.z, k2, eps, a.z, b.z, d.z);
    fm::grad_mu<ZINB>(xv.w, m.w, p.w, k3, eps, a.w, b.w, d.w);
    if (dmu) *reinterpret_cast<float4*>(dmu + r * G + g) = make_float4(go.x * a.x, go.y * a.y, go.z * a.z, go.w * a.w);
    if (ZINB && dpi) *reinterpret_cast<float4*>(dpi + r * G + g) = make_float4(go.x * d.x, go.y * d.y, go.z * d.z, go.w * d.w);
    acc.x = fmaf(go.x, b.x, acc.x);
    acc.y = fmaf(go.y, b.y, acc.y);
    acc.z = fmaf(go.z, b.z, acc.z);
    acc.w = fmaf(go.w, b.w, acc.w);
  }
  if (dtheta) *reinterpret_cast<float4*>(dtheta + (int64_t)blockIdx.y * G + g) = acc;  // per-row-tile partial
}

static inline bool lp_al16(const void* p) { return p == nullptr || ((uintptr_t)p % 16) == 0; }

// out[n] (+)= sum_t partials[t, n], fixed order
__global__ void colsum_partials_kernel(const float* __restrict__ p, int64_t T, int64_t N, float* __restrict__ out,
                                       int accumulate) {
  const int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  float s = 0.f;
  for (int64_t t = 0; t < T; ++t) s += p[t * N + n];
  out[n] = accumulate ? out[n] + s : s;
}

template <bool ZINB>
static int launch_fwd(const float* x, int64_t ldx, const float* mu, int64_t ldmu, const float* theta,
                      int64_t theta_ld, const float* pi, int64_t ldpi, int64_t B, int64_t G, float eps, float* out,
                      int64_t ldo, int32_t* bad, void* stream, const char* name) {
  SCVI_REQUIRE(x && mu && theta && out && (!ZINB || pi), "%s: NULL pointer", name);
  SCVI_REQUIRE(B >= 0 && G >= 0 && ldx >= G && ldmu >= G && ldo >= G, "%s: bad shape", name);
  if (B == 0 || G == 0) return SCVI_OK;
  dim3 grid((unsigned)cdiv(G, LP_THREADS), (unsigned)cdiv(B, SCVI_BN_TILE));
  SCVI_REQUIRE(grid.y <= 65535, "%s: B too large for one launch (max %d rows)", name, 65535 * SCVI_BN_TILE);
  if (theta_ld == 0 && G % 4 == 0 && ldx % 4 == 0 && ldmu % 4 == 0 && ldo % 4 == 0 && (!ZINB || ldpi % 4 == 0) && lp_al16(x) &&
      lp_al16(mu) && lp_al16(theta) && lp_al16(out) && lp_al16(pi)) {
    grid.x = (unsigned)cdiv(G / 4, LP4_THREADS);
    // few cells (totalVI / SCANVI minibatches): 16-row tiles, so that the launch still covers the SMs
    const int row_tile = ((int64_t)grid.x * grid.y >= 2 * 148) ? SCVI_BN_TILE : 16;
    grid.y = (unsigned)cdiv(B, row_tile);
    logprob_fwd4_kernel<ZINB><<<grid, LP4_THREADS, 0, (cudaStream_t)stream>>>(x, ldx, mu, ldmu, theta, pi, ldpi, B, G, eps, out, ldo, bad, row_tile);
    SCVI_LAUNCH_CHECK(name);
    return SCVI_OK;
  }
  if (theta_ld == 0)
    logprob_fwd_kernel<ZINB, true><<<grid, LP_THREADS, 0, (cudaStream_t)stream>>>(x, ldx, mu, ldmu, theta, 0, pi, ldpi, B, G, eps, out, ldo, bad);
  else
    logprob_fwd_kernel<ZINB, false><<<grid, LP_THREADS, 0, (cudaStream_t)stream>>>(x, ldx, mu, ldmu, theta, theta_ld, pi, ldpi, B, G, eps, out, ldo, bad);
  SCVI_LAUNCH_CHECK(name);
  return SCVI_OK;
}

template <bool ZINB>
static int launch_bwd(const float* x, int64_t ldx, const float* mu, int64_t ldmu, const float* theta,
                      int64_t theta_ld, const float* pi, int64_t ldpi, const float* gout, int64_t ldg, int64_t B,
                      int64_t G, float eps, float* dmu, float* dtheta, float* dpi, void* stream, const char* name) {
  SCVI_REQUIRE(x && mu && theta && gout && (!ZINB || pi), "%s: NULL pointer", name);
  SCVI_REQUIRE(B >= 0 && G >= 0 && ldx >= G && ldmu >= G && ldg >= G, "%s: bad shape", name);
  if (B == 0 || G == 0) return SCVI_OK;
  dim3 grid((unsigned)cdiv(G, LP_THREADS), (unsigned)cdiv(B, SCVI_BN_TILE));
  SCVI_REQUIRE(grid.y <= 65535, "%s: B too large for one launch", name);
  if (theta_ld == 0 && G % 4 == 0 && ldx % 4 == 0 && ldmu % 4 == 0 && ldg % 4 == 0 && (!ZINB || ldpi % 4 == 0) && lp_al16(x) &&
      lp_al16(mu) && lp_al16(theta) && lp_al16(gout) && lp_al16(pi) && lp_al16(dmu) && lp_al16(dtheta) && lp_al16(dpi)) {
    grid.x = (unsigned)cdiv(G / 4, LP4_THREADS);
    logprob_bwd4_kernel<ZINB><<<grid, LP4_THREADS, 0, (cudaStream_t)stream>>>(x, ldx, mu, ldmu, theta, pi, ldpi, gout, ldg, B, G, eps, dmu, dtheta, dpi);
    SCVI_LAUNCH_CHECK(name);
    return SCVI_OK;
  }
  if (theta_ld == 0)
    logprob_bwd_kernel<ZINB, true><<<grid, LP_THREADS, 0, (cudaStream_t)stream>>>(x, ldx, mu, ldmu, theta, 0, pi, ldpi, gout, ldg, B, G, eps, dmu, dtheta, dpi);
  else
    logprob_bwd_kernel<ZINB, false><<<grid, LP_THREADS, 0, (cudaStream_t)stream>>>(x, ldx, mu, ldmu, theta, theta_ld, pi, ldpi, gout, ldg, B, G, eps, dmu, dtheta, dpi);
  SCVI_LAUNCH_CHECK(name);
  return SCVI_OK;
}

}  // namespace scvi

extern "C" {

int scvi_nb_logprob_fwd(const float* x, int64_t ldx, const float* mu, int64_t ldmu, const float* theta,
                        int64_t theta_ld, int64_t B, int64_t G, float eps, float* out, int64_t ldo,
                        int32_t* bad_support, void* stream) {
  return scvi::launch_fwd<false>(x, ldx, mu, ldmu, theta, theta_ld, nullptr, 0, B, G, eps, out, ldo, bad_support, stream, "scvi_nb_logprob_fwd");
}
int scvi_zinb_logprob_fwd(const float* x, int64_t ldx, const float* mu, int64_t ldmu, const float* theta,
                          int64_t theta_ld, const float* pi, int64_t ldpi, int64_t B, int64_t G, float eps,
                          float* out, int64_t ldo, int32_t* bad_support, void* stream) {
  return scvi::launch_fwd<true>(x, ldx, mu, ldmu, theta, theta_ld, pi, ldpi, B, G, eps, out, ldo, bad_support, stream, "scvi_zinb_logprob_fwd");
}
int scvi_nb_logprob_bwd(const float* x, int64_t ldx, const float* mu, int64_t ldmu, const float* theta,
                        int64_t theta_ld, const float* gout, int64_t ldg, int64_t B, int64_t G, float eps,
                        float* dmu, float* dtheta, void* stream) {
  return scvi::launch_bwd<false>(x, ldx, mu, ldmu, theta, theta_ld, nullptr, 0, gout, ldg, B, G, eps, dmu, dtheta, nullptr, stream, "scvi_nb_logprob_bwd");
}
int scvi_zinb_logprob_bwd(const float* x, int64_t ldx, const float* mu, int64_t ldmu, const float* theta,
                          int64_t theta_ld, const float* pi, int64_t ldpi, const float* gout, int64_t ldg,
                          int64_t B, int64_t G, float eps, float* dmu, float* dtheta, float* dpi, void* stream) {
  return scvi::launch_bwd<true>(x, ldx, mu, ldmu, theta, theta_ld, pi, ldpi, gout, ldg, B, G, eps, dmu, dtheta, dpi, stream, "scvi_zinb_logprob_bwd");
}
int scvi_colsum_partials(const float* partials, int64_t n_tiles, int64_t N, float* out, int accumulate, void* stream) {
  using namespace scvi;
  SCVI_REQUIRE(partials && out && n_tiles >= 0 && N >= 0, "scvi_colsum_partials: bad argument");
  if (N == 0) return SCVI_OK;
  colsum_partials_kernel<<<(unsigned)cdiv(N, 256), 256, 0, (cudaStream_t)stream>>>(partials, n_tiles, N, out, accumulate);
  SCVI_LAUNCH_CHECK("scvi_colsum_partials");
  return SCVI_OK;
}
}
