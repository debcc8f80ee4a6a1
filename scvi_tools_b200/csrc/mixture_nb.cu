// Negative-binomial mixture log-likelihood (the TOTALVI protein likelihood), forward + backward, elementwise
// over [B,P].  Replaces log_mixture_nb with one shared inverse dispersion (theta_2 = None: the branch TOTALVI
// takes) as called by NegativeBinomialMixture.log_prob
// (reference src/scvi/distributions/_negative_binomial.py:138-216, :589-717):
//
//   l_k  = theta (log(theta+eps) - log(theta+mu_k+eps)) + x (log(mu_k+eps) - log(theta+mu_k+eps))
//          + lgamma(x+theta) - lgamma(theta) - lgamma(x+1)                         k = 1 (background), 2
//   out  = logsumexp(l_1, l_2 - pi) - softplus(-pi)                                pi = mixture logit
//
// The three lgamma terms and log(theta+eps) are shared by both components (computed once); the reference
// materialises ~20 [B,P] temporaries.  HBM-bound: forward reads x, mu_1, mu_2, pi (+theta) and writes out.
#include "likelihood.cuh"

namespace scvi {

constexpr int MX_THREADS = 128;
constexpr int MX_RG = 8;  // row groups per 64-row tile in the backward kernel (blockDim.y)

struct MixTerms {
  float l1, l2;  // component log-likelihoods
  float a;       // l_2 - pi
  float lse;     // logsumexp(l_1, a)
};

__device__ __forceinline__ MixTerms mix_terms(float x, float mu1, float mu2, float pi, const GeneConst& c, float eps) {
  const float lt1 = logf(c.theta + mu1 + eps), lt2 = logf(c.theta + mu2 + eps);
  const float shared = lgammaf(x + c.theta) - c.lg_theta - lgammaf(x + 1.f);
  MixTerms t;
  t.l1 = c.theta * (c.log_theta - lt1) + x * (logf(mu1 + eps) - lt1) + shared;
  t.l2 = c.theta * (c.log_theta - lt2) + x * (logf(mu2 + eps) - lt2) + shared;
  t.a = t.l2 - pi;
  const float m = fmaxf(t.l1, t.a);
  t.lse = (m == -INFINITY) ? -INFINITY : m + logf(expf(t.l1 - m) + expf(t.a - m));
  return t;
}

// thread = one protein column x a tile of SCVI_BN_TILE rows (per-protein constants hoisted when theta is [P])
template <bool THETA_PER_COL>
__global__ void __launch_bounds__(MX_THREADS)
    mixture_fwd_kernel(const float* __restrict__ x, int64_t ldx, const float* __restrict__ mu1, int64_t ld1,
                       const float* __restrict__ mu2, int64_t ld2, const float* __restrict__ theta, int64_t theta_ld,
                       const float* __restrict__ pi, int64_t ldpi, int64_t B, int64_t P, float eps,
                       float* __restrict__ out, int64_t ldo, int32_t* __restrict__ bad_support, int row_tile) {
  const int64_t g = (int64_t)blockIdx.x * MX_THREADS + threadIdx.x;
  const int64_t r0 = (int64_t)blockIdx.y * row_tile;
  const int64_t r1 = min(r0 + (int64_t)row_tile, B);
  if (g >= P) return;
  GeneConst c;
  if (THETA_PER_COL) c = make_gene_const(theta[g], eps);
  bool bad = false;
  for (int64_t r = r0; r < r1; ++r) {
    const float xv = x[r * ldx + g];
    if (!THETA_PER_COL) c = make_gene_const(theta[r * theta_ld + g], eps);
    const float p = pi[r * ldpi + g];
    const MixTerms t = mix_terms(xv, mu1[r * ld1 + g], mu2[r * ld2 + g], p, c, eps);
    out[r * ldo + g] = t.lse - softplus_t(-p);
    bad |= (xv < 0.f) || (xv != floorf(xv));
  }
  if (bad_support != nullptr && bad) *bad_support = 1;
}

template <bool THETA_PER_COL>
__global__ void __launch_bounds__(MX_THREADS * MX_RG)
    mixture_bwd_kernel(const float* __restrict__ x, int64_t ldx, const float* __restrict__ mu1, int64_t ld1,
                       const float* __restrict__ mu2, int64_t ld2, const float* __restrict__ theta, int64_t theta_ld,
                       const float* __restrict__ pi, int64_t ldpi, const float* __restrict__ gout, int64_t ldg,
                       int64_t B, int64_t P, float eps, float* __restrict__ dmu1, float* __restrict__ dmu2,
                       float* __restrict__ dtheta, float* __restrict__ dpi) {
  // blockDim = (MX_THREADS, MX_RG): row group ty owns SCVI_BN_TILE / MX_RG consecutive rows of the tile; the
  // per-tile dtheta partial is the sum of the groups' partials in group order (fixed)
  __shared__ float dth_sh[MX_RG][MX_THREADS];
  const int64_t g = (int64_t)blockIdx.x * MX_THREADS + threadIdx.x;
  const int64_t t0 = (int64_t)blockIdx.y * SCVI_BN_TILE;
  const int64_t r0 = min(t0 + (int64_t)threadIdx.y * (SCVI_BN_TILE / MX_RG), B);
  const int64_t r1 = min(r0 + (int64_t)(SCVI_BN_TILE / MX_RG), B);
  const bool col_ok = g < P;
  GeneConst c;
  if (THETA_PER_COL && col_ok) c = make_gene_const(theta[g], eps);
  float dth_acc = 0.f;
  for (int64_t r = r0; col_ok && r < r1; ++r) {
    const float xv = x[r * ldx + g];
    const float m1 = mu1[r * ld1 + g], m2 = mu2[r * ld2 + g];
    if (!THETA_PER_COL) c = make_gene_const(theta[r * theta_ld + g], eps);
    const float p = pi[r * ldpi + g];
    const float go = gout[r * ldg + g];
    const MixTerms t = mix_terms(xv, m1, m2, p, c, eps);
    // responsibilities of the two components
    const float w1 = (t.lse == -INFINITY) ? 0.5f : expf(t.l1 - t.lse);
    const float w2 = (t.lse == -INFINITY) ? 0.5f : expf(t.a - t.lse);
    float d1, dt1, d2, dt2;
    nb_ll_grad(xv, m1, c, eps, d1, dt1);
    nb_ll_grad(xv, m2, c, eps, d2, dt2);
    if (dmu1) dmu1[r * P + g] = go * w1 * d1;
    if (dmu2) dmu2[r * P + g] = go * w2 * d2;
    if (dpi) dpi[r * P + g] = go * (sigmoid_t(-p) - w2);
    if (dtheta) {
      const float v = go * (w1 * dt1 + w2 * dt2);
      if (THETA_PER_COL)
        dth_acc += v;
      else
        dtheta[r * P + g] = v;
    }
  }
  if (THETA_PER_COL && dtheta) {  // per-row-tile partial
    dth_sh[threadIdx.y][threadIdx.x] = dth_acc;
    __syncthreads();
    if (threadIdx.y == 0 && col_ok) {
      float t = 0.f;
#pragma unroll
      for (int q = 0; q < MX_RG; ++q) t += dth_sh[q][threadIdx.x];
      dtheta[(int64_t)blockIdx.y * P + g] = t;
    }
  }
}

}  // namespace scvi

extern "C" {

int scvi_mixture_nb_logprob_fwd(const float* x, int64_t ldx, const float* mu1, int64_t ldmu1, const float* mu2,
                                int64_t ldmu2, const float* theta, int64_t theta_ld, const float* pi, int64_t ldpi,
                                int64_t B, int64_t P, float eps, float* out, int64_t ldo, int32_t* bad_support,
                                void* stream) {
  using namespace scvi;
  SCVI_REQUIRE(x && mu1 && mu2 && theta && pi && out, "scvi_mixture_nb_logprob_fwd: NULL pointer");
  SCVI_REQUIRE(B >= 0 && P >= 0 && ldx >= P && ldmu1 >= P && ldmu2 >= P && ldpi >= P && ldo >= P &&
                   (theta_ld == 0 || theta_ld >= P),
               "scvi_mixture_nb_logprob_fwd: bad shape");
  if (B == 0 || P == 0) return SCVI_OK;
  // a thread walks `row_tile` rows of its column: short tiles when 64-row tiles would leave most SMs without work
  // (totalVI: 256 cells x ~100 proteins)
  const int gx = (int)cdiv(P, MX_THREADS);
  const int row_tile = (gx * cdiv(B, SCVI_BN_TILE) >= 148) ? SCVI_BN_TILE : 4;
  dim3 grid((unsigned)gx, (unsigned)cdiv(B, row_tile));
  SCVI_REQUIRE(grid.y <= 65535, "scvi_mixture_nb_logprob_fwd: B too large for one launch");
  cudaStream_t st = (cudaStream_t)stream;
  if (theta_ld == 0)
    mixture_fwd_kernel<true><<<grid, MX_THREADS, 0, st>>>(x, ldx, mu1, ldmu1, mu2, ldmu2, theta, 0, pi, ldpi, B, P, eps, out, ldo, bad_support, row_tile);
  else
    mixture_fwd_kernel<false><<<grid, MX_THREADS, 0, st>>>(x, ldx, mu1, ldmu1, mu2, ldmu2, theta, theta_ld, pi, ldpi, B, P, eps, out, ldo, bad_support, row_tile);
  SCVI_LAUNCH_CHECK("scvi_mixture_nb_logprob_fwd");
  return SCVI_OK;
}

int scvi_mixture_nb_logprob_bwd(const float* x, int64_t ldx, const float* mu1, int64_t ldmu1, const float* mu2,
                                int64_t ldmu2, const float* theta, int64_t theta_ld, const float* pi, int64_t ldpi,
                                const float* gout, int64_t ldg, int64_t B, int64_t P, float eps, float* dmu1,
                                float* dmu2, float* dtheta, float* dpi, void* stream) {
  using namespace scvi;
  SCVI_REQUIRE(x && mu1 && mu2 && theta && pi && gout, "scvi_mixture_nb_logprob_bwd: NULL pointer");
  SCVI_REQUIRE(B >= 0 && P >= 0 && ldx >= P && ldmu1 >= P && ldmu2 >= P && ldpi >= P && ldg >= P &&
                   (theta_ld == 0 || theta_ld >= P),
               "scvi_mixture_nb_logprob_bwd: bad shape");
  if (B == 0 || P == 0) return SCVI_OK;
  dim3 grid((unsigned)cdiv(P, MX_THREADS), (unsigned)cdiv(B, SCVI_BN_TILE));
  cudaStream_t st = (cudaStream_t)stream;
  const dim3 block(MX_THREADS, MX_RG);
  if (theta_ld == 0)
    mixture_bwd_kernel<true><<<grid, block, 0, st>>>(x, ldx, mu1, ldmu1, mu2, ldmu2, theta, 0, pi, ldpi, gout, ldg, B, P, eps, dmu1, dmu2, dtheta, dpi);
  else
    mixture_bwd_kernel<false><<<grid, block, 0, st>>>(x, ldx, mu1, ldmu1, mu2, ldmu2, theta, theta_ld, pi, ldpi, gout, ldg, B, P, eps, dmu1, dmu2, dtheta, dpi);
  SCVI_LAUNCH_CHECK("scvi_mixture_nb_logprob_bwd");
  return SCVI_OK;
}
}
