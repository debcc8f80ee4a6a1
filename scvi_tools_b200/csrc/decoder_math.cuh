// Per-element math of the fused decoder kernels (shared by the FFMA and the tcgen05 variants).
#pragma once
#include "likelihood.cuh"

namespace scvi {

struct RowC {  // per-row constants
  float m, invl, c, S, u;
};

// ll and d(ll)/d(mu) of one element, sharing sub-expressions
template <bool ZINB>
__device__ __forceinline__ void elem_fwd(float x, float mu, float pi, const GeneConst& gc, float eps, float& ll,
                                         float& dmu) {
  const float T = gc.theta + mu + eps;
  const float rT = 1.f / T;
  const float d_log = gc.log_theta - logf(T);
  if (ZINB) {
    const float sp = softplus_t(-pi);
    const float a = -pi + gc.theta * d_log;
    if (x < eps) {
      ll = softplus_t(a) - sp;
      dmu = sigmoid_t(a) * (-gc.theta * rT);
    } else if (x > eps) {
      const float me = mu + eps;
      ll = -sp + a + x * (logf(me) - logf(T)) + lgammaf(x + gc.theta) - gc.lg_theta - lgammaf(x + 1.f);
      dmu = -gc.theta * rT + x * (1.f / me - rT);
    } else {
      ll = 0.f;
      dmu = 0.f;
    }
  } else {
    const float me = mu + eps;
    ll = gc.theta * d_log;
    dmu = -gc.theta * rT;
    if (x != 0.f) {  // x == 0: the remaining terms are exactly 0 in the reference too
      ll += x * (logf(me) - logf(T)) + lgammaf(x + gc.theta) - gc.lg_theta - lgammaf(x + 1.f);
      dmu += x * (1.f / me - rT);
    }
  }
}

// per-element backward: ds, dp (already multiplied by the upstream u) and u * theta * d(reconst)/d(theta)
template <bool ZINB>
__device__ __forceinline__ void elem_bwd(float x, float s, float pi, const RowC& rc, const GeneConst& gc, float eps,
                                         float& ds, float& dp, float& dr) {
  const float e = expf(s - rc.m);
  const float mu = e * rc.c;
  float dmu, dth, dpi = 0.f;
  if (ZINB)
    zinb_ll_grad(x, mu, pi, gc, eps, dmu, dth, dpi);
  else
    nb_ll_grad(x, mu, gc, eps, dmu, dth);
  // reconst = -sum ll:  d/ds = mu * g - softmax * S,  g = -dmu,  softmax = e * invl
  ds = rc.u * (-mu * dmu - e * rc.invl * rc.S);
  dp = -rc.u * dpi;
  dr = -rc.u * gc.theta * dth;
}

}  // namespace scvi
