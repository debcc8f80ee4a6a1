// Per-element NB / ZINB log-likelihood math shared by the standalone and the fused kernels.
// Follows log_nb_positive / log_zinb_positive
// (reference src/scvi/distributions/_negative_binomial.py:99-135, :36-96), fp32.
#pragma once
#include "common.cuh"

namespace scvi {

// F.softplus(x) with torch defaults beta=1, threshold=20
__device__ __forceinline__ float softplus_t(float x) { return x > 20.f ? x : log1pf(expf(x)); }
__device__ __forceinline__ float sigmoid_t(float x) { return 1.f / (1.f + expf(-x)); }

// digamma for x > 0 (recurrence to x >= 6, then the asymptotic series)
__device__ __forceinline__ float digamma_pos(float x) {
  float acc = 0.f;
  while (x < 6.f) {
    acc -= 1.f / x;
    x += 1.f;
  }
  const float r = 1.f / x, r2 = r * r;
  // 1/12 - r2/120 + r4/252 - r6/240 + r8/132
  float s = r2 * (0.083333333333f - r2 * (0.0083333333333f - r2 * (0.003968253968f - r2 * (0.004166666667f - r2 * 0.007575757576f))));
  return acc + logf(x) - 0.5f * r - s;
}

// psi(x + theta) - psi(theta) for integer-valued x > 0: exact finite sum for small counts
__device__ __forceinline__ float digamma_diff(float x, float theta) {
  if (x <= 12.f) {
    float s = 0.f;
    for (float k = 0.f; k < x; k += 1.f) s += 1.f / (theta + k);
    return s;
  }
  return digamma_pos(x + theta) - digamma_pos(theta);
}

// per-gene constants hoisted out of the per-element work
struct GeneConst {
  float theta;      // exp(px_r)
  float log_theta;  // log(theta + eps)
  float lg_theta;   // lgamma(theta)
};
__device__ __forceinline__ GeneConst make_gene_const(float theta, float eps) {
  GeneConst c;
  c.theta = theta;
  c.log_theta = logf(theta + eps);
  c.lg_theta = lgammaf(theta);
  return c;
}

// NB log-likelihood of one element
__device__ __forceinline__ float nb_ll(float x, float mu, const GeneConst& c, float eps) {
  const float l_tm = logf(c.theta + mu + eps);
  return c.theta * (c.log_theta - l_tm) + x * (logf(mu + eps) - l_tm) + lgammaf(x + c.theta) - c.lg_theta -
         lgammaf(x + 1.f);
}

// ZINB log-likelihood of one element (pi = zero-inflation logit)
__device__ __forceinline__ float zinb_ll(float x, float mu, float pi, const GeneConst& c, float eps) {
  const float sp = softplus_t(-pi);
  const float l_tm = logf(c.theta + mu + eps);
  const float a = -pi + c.theta * (c.log_theta - l_tm);
  if (x < eps) return softplus_t(a) - sp;
  if (x > eps)
    return -sp + a + x * (logf(mu + eps) - l_tm) + lgammaf(x + c.theta) - c.lg_theta - lgammaf(x + 1.f);
  return 0.f;  // x == eps: both reference masks are 0
}

// d(ll)/d(mu), d(ll)/d(theta) for NB
__device__ __forceinline__ void nb_ll_grad(float x, float mu, const GeneConst& c, float eps, float& dmu,
                                           float& dtheta) {
  const float T = c.theta + mu + eps;
  const float rT = 1.f / T;
  dmu = -c.theta * rT + x * (1.f / (mu + eps) - rT);
  dtheta = (c.log_theta - logf(T)) + c.theta * (1.f / (c.theta + eps) - rT) - x * rT +
           (x > 0.f ? digamma_diff(x, c.theta) : 0.f);
}

// ZINB: returns ll and the three partials in one go
__device__ __forceinline__ void zinb_ll_grad(float x, float mu, float pi, const GeneConst& c, float eps,
                                             float& dmu, float& dtheta, float& dpi) {
  const float T = c.theta + mu + eps;
  const float rT = 1.f / T;
  const float d_log = c.log_theta - logf(T);
  const float da_dtheta = d_log + c.theta * (1.f / (c.theta + eps) - rT);
  if (x < eps) {
    const float a = -pi + c.theta * d_log;
    const float sa = sigmoid_t(a);
    dmu = sa * (-c.theta * rT);
    dtheta = sa * da_dtheta;
    dpi = sigmoid_t(-pi) - sa;
  } else if (x > eps) {
    dmu = -c.theta * rT + x * (1.f / (mu + eps) - rT);
    dtheta = da_dtheta - x * rT + digamma_diff(x, c.theta);
    dpi = -sigmoid_t(pi);
  } else {
    dmu = dtheta = dpi = 0.f;
  }
}

}  // namespace scvi
