// Hand-written fp32 per-element math of the fused decoder kernels: NB / ZINB log-likelihood and its
// partial derivatives on the SFU approximations (ex2 / lg2 / rcp, ~1-2 ulp) instead of libdevice
// lgammaf / log1pf / expf.  Follows log_nb_positive / log_zinb_positive
// (reference src/scvi/distributions/_negative_binomial.py:99-135, :36-96) term by term; the
// derivatives are those autograd produces for the same expressions.  Accuracy budget: per-element
// absolute error ~1e-6 (see tests/test_fastmath_host.py, which compiles this header for the host and
// compares with float64), far inside the 1e-5 relative per-cell tolerance.
//
// The header also compiles as plain host C++ (SCVI_FASTMATH_HOST) with exp2f/log2f standing in for
// the SFU instructions, so the formulas are unit-tested without a GPU.
#pragma once
#include <math.h>
#include <stdint.h>

#ifdef __CUDACC__
#define SCVI_FM __device__ __forceinline__
#else
#define SCVI_FM static inline
#endif

#include <string.h>

namespace scvi {
namespace fm {

constexpr float LOG2E = 1.4426950408889634f;
constexpr float LN2 = 0.6931471805599453f;
constexpr float HALF_LOG_2PI = 0.9189385332046727f;
constexpr int LGX1_TABLE = 256;  // lgamma(x + 1) table for integer counts below this

#if defined(__CUDA_ARCH__)
SCVI_FM float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
SCVI_FM float lg2(float x) {
  float y;
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
SCVI_FM float rcp(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
#else
SCVI_FM float ex2(float x) { return exp2f(x); }
SCVI_FM float lg2(float x) { return log2f(x); }
SCVI_FM float rcp(float x) { return 1.0f / x; }
#endif

// 2^x for x <= 0 on the FMA / ALU pipes only (no SFU op): round-to-nearest split x = i + f with the 1.5 * 2^23 trick,
// degree-5 minimax polynomial of 2^f on [-0.5, 0.5] (relative error 2.2e-7 in fp32 Horner form: the accuracy of
// ex2.approx), exponent patched in with an integer add.  The fused likelihood kernels are bound by the SFU pipe
// inside their element math (9 MUFU per ZINB element = 72 of ~118 issue cycles per warp and element) while the FMA
// pipe idles at 19 %.  MEASURED (r02): using this for the two exp(-|.|) of the zero-inflation terms (ZINB: 9 -> 7 MUFU
// per element, +22 FMA / ALU instructions) made K2 119.8 us (118.6) and K3 104.3 us (99.6): the element math is bound
// by issue slots and dependent latency as much as by the SFU pipe, so trading 1 MUFU for 11 other instructions loses.
// Kept (with its accuracy test) for kernels whose SFU pipe is the sole bound; not used by the fused kernels.
SCVI_FM float ex2_fma(float x) {
  x = fmaxf(x, -125.f);
  const float t = x + 12582912.f;
  const float f = x - (t - 12582912.f);
  float p = 0.0013266970386324052f;
  p = fmaf(p, f, 0.009675459745517628f);
  p = fmaf(p, f, 0.05550742616002558f);
  p = fmaf(p, f, 0.24022121753561645f);
  p = fmaf(p, f, 0.6931469491610631f);
  p = fmaf(p, f, 1.0000000710296983f);
#if defined(__CUDA_ARCH__)
  return __int_as_float(__float_as_int(p) + (__float_as_int(t) << 23));
#else
  int pb, tb;
  memcpy(&pb, &p, 4);
  memcpy(&tb, &t, 4);
  pb += (int)((unsigned)tb << 23);
  float r;
  memcpy(&r, &pb, 4);
  return r;
#endif
}

SCVI_FM int fbits(float x) {
#if defined(__CUDA_ARCH__)
  return __float_as_int(x);
#else
  int b;
  memcpy(&b, &x, 4);
  return b;
#endif
}

// Is x a non-negative integer below `limit` (<= 2^22)?  If so idx = x.  No float<->int conversion instruction:
// F2I / I2F / FRND issue on the same quarter-rate pipe as the transcendentals, which bounds these kernels; x + 2^23
// carries round(x) in its low mantissa bits and (x + 2^23) - 2^23 == x exactly when x is an integer below 2^23.
SCVI_FM bool count_index(float x, int limit, int& idx) {
  const float t = x + 8388608.f;
  idx = fbits(t) & 0x7fffff;
  return (t - 8388608.f == x) & (x >= 0.f) & (x < (float)limit);
}

// support check of the count distributions: negative, NaN or non-integer (floats from 2^23 up are all integers)
SCVI_FM bool bad_count(float x) {
  const float t = x + 8388608.f;
  return !(x >= 0.f) | ((x < 8388608.f) & (t - 8388608.f != x));
}

SCVI_FM float fexp(float x) { return ex2(x * LOG2E); }
SCVI_FM float flog(float x) { return lg2(x) * LN2; }

// Stirling series of lgamma at w >= 4 given ln(w) and 1/w (truncation error < 4e-8 at w = 4)
SCVI_FM float stirling(float w, float lw, float rw) {
  const float rw2 = rw * rw;
  const float ser = rw * (0.083333333333f + rw2 * (-0.0027777777778f + rw2 * (0.00079365079365f - rw2 * 0.00059523809524f)));
  return fmaf(w - 0.5f, lw, -w) + HALF_LOG_2PI + ser;
}

// lgamma(z), z > 0:  lgamma(z) = stirling(z + 4) - ln(z (z+1) (z+2) (z+3)).   3 SFU ops.
SCVI_FM float lgamma_pos(float z) {
  const float w = z + 4.f;
  const float q = fmaf(z, z, 3.f * z);  // z (z + 3)
  const float P = q * (q + 2.f);        // z (z+1) (z+2) (z+3)
  return stirling(w, flog(w), rcp(w)) - flog(P);
}

// digamma(z), z > 0:  digamma(z + 4) - sum_{k<4} 1/(z + k);  the sum is P'(z) / P(z).   3 SFU ops.
SCVI_FM float digamma_pos_fast(float z) {
  const float w = z + 4.f;
  const float q = fmaf(z, z, 3.f * z);
  const float P = q * (q + 2.f);
  const float dP = (2.f * q + 2.f) * (2.f * z + 3.f);
  const float rw = rcp(w), rw2 = rw * rw;
  const float ser = rw2 * (0.083333333333f - rw2 * (0.0083333333333f - rw2 * (0.003968253968f - rw2 * 0.004166666667f)));
  return flog(w) - 0.5f * rw - ser - dP * rcp(P);
}

// lgamma(x + 1) for a count x >= 0: table for integer x < LGX1_TABLE, series otherwise
SCVI_FM float lgamma_x1(float x, const float* __restrict__ table) {
  int xi;
  if (count_index(x, LGX1_TABLE, xi)) return table[xi];
  return lgamma_pos(x + 1.f);
}

// per-gene constants hoisted out of the per-element work (computed with libdevice precision)
struct GeneK {
  float theta;      // exp(px_r)
  float log_theta;  // log(theta + eps)
  float lg_theta;   // lgamma(theta)
  float rth;        // 1 / (theta + eps)
  float dg_theta;   // digamma(theta)
};

// One element of the fused forward (+ the pieces of the backward that depend only on this element).
//   s, pi : scale / dropout logits (bias included);  krow = -m*log2(e) + log2(exp(library)/sumexp)
//   out: ll = log-likelihood, g1 = -mu * d(ll)/d(mu)  (so d(reconst)/d(s) = g1 - softmax * sum_g g1),
//        dpi = d(ll)/d(pi), mu
template <bool ZINB>
SCVI_FM void elem_fwd(float x, float s, float pi, float krow, const GeneK& g, float eps,
                      const float* __restrict__ lgx1, float& ll, float& g1, float& dpi, float& mu) {
  mu = ex2(fmaf(s, LOG2E, krow));
  const float T = g.theta + mu + eps;
  const float rT = rcp(T);
  const float me = mu + eps;
  // log(theta+eps) - log(theta+mu+eps) = -log(1 + mu/(theta+eps)): one log, no cancellation for mu << theta
  const float d_log = -flog(fmaf(mu, g.rth, 1.f));
  if (ZINB) {
    const float a = fmaf(g.theta, d_log, -pi);
    const float ep = ex2(-fabsf(pi) * LOG2E);  // exp(-|pi|)
    const float rp = rcp(1.f + ep);            // sigmoid(|pi|)
    const float ea = ex2(-fabsf(a) * LOG2E);
    const bool zero = x < eps;
    const float r1 = rcp(zero ? 1.f + ea : me);
    if (zero) {
      // softplus(a) - softplus(-pi) = max(a,0) - max(-pi,0) + ln((1+ea) * sigmoid(|pi|))
      ll = fmaxf(a, 0.f) - fmaxf(-pi, 0.f) + flog((1.f + ea) * rp);
      const float sa = a >= 0.f ? r1 : ea * r1;  // sigmoid(a)
      g1 = sa * g.theta * mu * rT;
      dpi = (pi >= 0.f ? ep * rp : rp) - sa;  // sigmoid(-pi) - sigmoid(a)
    } else if (x > eps) {
      const float z = x + g.theta;
      const float w = z + 4.f;
      const float q = fmaf(z, z, 3.f * z);
      const float P = q * (q + 2.f);
      // -softplus(-pi) - ln P = -max(-pi,0) - ln((1+ep) * P)
      ll = a - fmaxf(-pi, 0.f) - flog((1.f + ep) * P) + x * flog(me * rT) + stirling(w, flog(w), rcp(w)) -
           g.lg_theta - lgamma_x1(x, lgx1);
      g1 = g.theta * mu * rT * (1.f - x * r1);  // r1 - rT = theta / (me T)
      dpi = -(pi >= 0.f ? rp : ep * rp);  // -sigmoid(pi)
    } else {  // x == eps: both reference masks are 0
      ll = 0.f;
      g1 = 0.f;
      dpi = 0.f;
    }
  } else {
    ll = g.theta * d_log;
    g1 = g.theta * mu * rT;
    dpi = 0.f;
    if (x != 0.f) {  // x == 0: the remaining terms are exactly 0 in the reference too
      ll += x * flog(me * rT) + lgamma_pos(x + g.theta) - g.lg_theta - lgamma_x1(x, lgx1);
      g1 *= 1.f - x * rcp(me);
    }
  }
}

// ---- table-driven, branch-free variant for the fused training kernel --------------------------------------
// Counts are small integers, so the only per-element lgamma work of the likelihood,
//     lgd(x, theta) = lgamma(x + theta) - lgamma(theta) - lgamma(x + 1),
// is read from a per-gene table of LGT entries (built per step from the current theta in double precision;
// lgd(0) = 0 exactly); x >= LGT or non-integer x falls back to the series.  The x == 0 and x > 0 cases of the
// zero-inflated likelihood share every SFU op through selected arguments (no divergent branches):
// 9 SFU ops per element (ZINB), 5 (NB).
constexpr int LGT = 64;

template <bool ZINB>
SCVI_FM void elem_fwd_tab(float x, float s, float pi, float krow, const GeneK& g, float eps,
                          const float* __restrict__ tab /*[LGT] of this gene*/, const float* __restrict__ lgx1,
                          float& ll, float& g1, float& dpi, float& mu) {
  mu = ex2(fmaf(s, LOG2E, krow));
  const float me = mu + eps;
  const float rT = rcp(g.theta + mu + eps);
  const float d_log = -flog(fmaf(mu, g.rth, 1.f));  // log(theta+eps) - log(theta+mu+eps)
  int xi;
  float lgd;
  if (count_index(x, LGT, xi)) lgd = tab[xi];
  else lgd = lgamma_pos(x + g.theta) - g.lg_theta - lgamma_x1(x, lgx1);  // rare
  const float base = g.theta * mu * rT;
  if (ZINB) {
    const bool zero = x < eps;
    const float a = fmaf(g.theta, d_log, -pi);
    const float ep = ex2(-fabsf(pi) * LOG2E);  // exp(-|pi|)
    const float op = 1.f + ep;
    const float rp = rcp(op);                   // sigmoid(|pi|)
    const float spn = fmaxf(-pi, 0.f) + flog(op);  // softplus(-pi)
    const float ea = ex2(-fabsf(a) * LOG2E);
    const float oa = 1.f + ea;
    const float r1 = rcp(zero ? oa : me);
    const float lsel = flog(zero ? oa : me * rT);  // zero: softplus(a) - max(a,0);  else log(mu+eps) - log(theta+mu+eps)
    const float sa = a >= 0.f ? r1 : ea * r1;      // sigmoid(a) when zero
    ll = (zero ? fmaxf(a, 0.f) : a) - spn + (zero ? 1.f : x) * lsel + lgd;
    g1 = base * (zero ? sa : fmaf(-x, r1, 1.f));   // r1 - rT = theta / (me T)
    const float e_rp = ep * rp;
    dpi = zero ? (pi >= 0.f ? e_rp : rp) - sa : -(pi >= 0.f ? rp : e_rp);
  } else {
    ll = fmaf(g.theta, d_log, fmaf(x, flog(me * rT), lgd));
    g1 = base * fmaf(-x, rcp(me), 1.f);
    dpi = 0.f;
  }
}

// ---- lean straight-line variants for the fused training kernels ------------------------------------------
// Contract: x is a non-negative INTEGER count below the table length (the caller checks its whole 8-element
// chunk once and takes the generic variant above otherwise), `lgd` / `dgd` are the table entries of (gene, x).
// No data-dependent branch, no float<->int conversion (both run on the SFU pipe): eight of these unrolled side
// by side interleave freely, which is what hides the SFU latency.  log(theta+mu+eps) is formed as
// log(theta+eps) + log(1 + mu/(theta+eps)) from the one lg2 the NB term needs anyway.
// SFU ops: ZINB 9 (ex2 x3, lg2 x3, rcp x3), NB 5.
template <bool ZINB>
SCVI_FM void elem_fwd_lean(float x, float lgd, float s, float pi, float krow, const GeneK& g, float eps, float& ll,
                           float& g1, float& dpi, float& mu) {
  mu = ex2(fmaf(s, LOG2E, krow));
  const float me = mu + eps;
  const float rT = rcp(g.theta + me);
  const float lu = lg2(fmaf(mu, g.rth, 1.f));        // log2((theta+mu+eps) / (theta+eps))
  const float thl = g.theta * (-LN2);
  const float base = g.theta * mu * rT;
  if (ZINB) {
    const bool zero = x == 0.f;
    const float a = fmaf(thl, lu, -pi);              // theta * (log(theta+eps) - log(theta+mu+eps)) - pi
    const float ep = ex2(-fabsf(pi) * LOG2E);        // exp(-|pi|)
    const float op = 1.f + ep;
    const float ea = ex2(-fabsf(a) * LOG2E);
    const float sel = zero ? 1.f + ea : me;
    const float rp = rcp(op);                        // sigmoid(|pi|)
    const float r1 = rcp(sel);
    const float l1 = lg2(op);                        // softplus(-pi) = max(-pi, 0) + ln2 * l1
    const float l2 = lg2(sel);                       // zero: softplus(a) - max(a, 0);  else log2(mu + eps)
    const float nz = fmaf(x, fmaf(LN2, l2 - lu, -g.log_theta), lgd);  // x (log(mu+eps) - log(theta+mu+eps)) + lgd
    const float head = (zero ? fmaxf(a, 0.f) : a) - fmaxf(-pi, 0.f);
    ll = fmaf(-LN2, l1, head) + (zero ? LN2 * l2 : nz);
    const float sa = a >= 0.f ? r1 : ea * r1;        // sigmoid(a) when zero
    g1 = base * (zero ? sa : fmaf(-x, r1, 1.f));
    const float e_rp = ep * rp;
    const float sp = pi >= 0.f ? e_rp : rp;          // sigmoid(-pi)
    const float sn = pi >= 0.f ? rp : e_rp;          // sigmoid(pi)
    dpi = zero ? sp - sa : -sn;
  } else {
    const float l2 = lg2(me);
    ll = fmaf(thl, lu, fmaf(x, fmaf(LN2, l2 - lu, -g.log_theta), lgd));
    g1 = base * fmaf(-x, rcp(me), 1.f);
    dpi = 0.f;
  }
}

// backward twin (genes-owned dW kernel): g1, dpi, dth = d(ll)/d(theta), mu.  SFU ops: ZINB 7, NB 4.
template <bool ZINB>
SCVI_FM void elem_bwd_lean(float x, float dgd, float s, float pi, float krow, const GeneK& g, float eps, float& g1,
                           float& dpi, float& dth, float& mu) {
  mu = ex2(fmaf(s, LOG2E, krow));
  const float me = mu + eps;
  const float rT = rcp(g.theta + me);
  const float lu = lg2(fmaf(mu, g.rth, 1.f));
  const float d_log = -LN2 * lu;
  const float da_dth = fmaf(g.theta, g.rth - rT, d_log);
  const float base = g.theta * mu * rT;
  const float nz_dth = fmaf(-x, rT, da_dth) + dgd;
  if (ZINB) {
    const bool zero = x == 0.f;
    const float a = fmaf(g.theta, d_log, -pi);
    const float ep = ex2(-fabsf(pi) * LOG2E);
    const float ea = ex2(-fabsf(a) * LOG2E);
    const float rp = rcp(1.f + ep);
    const float r1 = rcp(zero ? 1.f + ea : me);
    const float sa = a >= 0.f ? r1 : ea * r1;
    const float e_rp = ep * rp;
    g1 = base * (zero ? sa : fmaf(-x, r1, 1.f));
    dpi = zero ? (pi >= 0.f ? e_rp : rp) - sa : -(pi >= 0.f ? rp : e_rp);
    dth = zero ? sa * da_dth : nz_dth;
  } else {
    g1 = base * fmaf(-x, rcp(me), 1.f);
    dpi = 0.f;
    dth = nz_dth;  // x == 0: dgd = 0 and x * rT = 0
  }
}

// ---- the same likelihoods with mu given (standalone log_prob kernels) -------------------------------------
// lgamma(x + theta) - lgamma(theta) - lgamma(x + 1); exactly 0 at x == 0 as in the reference
SCVI_FM float lgd_series(float x, const GeneK& g, const float* __restrict__ lgx1) {
  // evaluated for x == 0 too and then discarded: no divergent branch around the three SFU ops (a warp of 128 elements
  // is never all-zero in practice, so the branch saved nothing and kept the elements from interleaving)
  const float v = lgamma_pos(x + g.theta) - g.lg_theta - lgamma_x1(x, lgx1);
  return x == 0.f ? 0.f : v;
}

// The same value as ll_mu below for a count that passed count_index(x, LGX1_TABLE): the lgamma(x + 1) table is read
// unconditionally, so the function is straight-line code -- several of them unrolled side by side interleave their
// SFU chains freely (a per-element fallback branch pins every element between two convergence barriers).
template <bool ZINB>
SCVI_FM float ll_mu_count(float x, float mu, float pi, const GeneK& g, float eps, const float* __restrict__ lgx1) {
  const float me = mu + eps;
  const float rT = rcp(g.theta + mu + eps);
  const float d_log = -flog(fmaf(mu, g.rth, 1.f));  // log(theta+eps) - log(theta+mu+eps)
  const float lx1 = lgx1[fbits(x + 8388608.f) & 0x7fffff];
  const float v = lgamma_pos(x + g.theta) - g.lg_theta - lx1;
  const float lgd = x == 0.f ? 0.f : v;
  if (!ZINB) return fmaf(g.theta, d_log, fmaf(x, flog(me * rT), lgd));
  const bool zero = x < eps;
  const float a = fmaf(g.theta, d_log, -pi);
  const float spn = fmaxf(-pi, 0.f) + flog(1.f + ex2(-fabsf(pi) * LOG2E));  // softplus(-pi)
  const float oa = 1.f + ex2(-fabsf(a) * LOG2E);
  const float lsel = flog(zero ? oa : me * rT);
  const float ll = (zero ? fmaxf(a, 0.f) : a) - spn + (zero ? 1.f : x) * lsel + lgd;
  return x == eps ? 0.f : ll;  // x == eps: both reference masks are 0
}

// all four counts of a 16-byte piece are integers in [0, LGX1_TABLE): the precondition of ll_mu_count
SCVI_FM bool counts_in_table4(float x0, float x1, float x2, float x3) {
  const float h = 0.5f * (float)(LGX1_TABLE - 1);
  bool ok = (x0 + 8388608.f) - 8388608.f == x0;
  ok &= (x1 + 8388608.f) - 8388608.f == x1;
  ok &= (x2 + 8388608.f) - 8388608.f == x2;
  ok &= (x3 + 8388608.f) - 8388608.f == x3;
  ok &= fabsf(x0 - h) <= h;
  ok &= fabsf(x1 - h) <= h;
  ok &= fabsf(x2 - h) <= h;
  ok &= fabsf(x3 - h) <= h;
  return ok;
}

template <bool ZINB>
SCVI_FM float ll_mu(float x, float mu, float pi, const GeneK& g, float eps, const float* __restrict__ lgx1) {
  const float me = mu + eps;
  const float rT = rcp(g.theta + mu + eps);
  const float d_log = -flog(fmaf(mu, g.rth, 1.f));  // log(theta+eps) - log(theta+mu+eps)
  const float lgd = lgd_series(x, g, lgx1);
  if (!ZINB) return fmaf(g.theta, d_log, fmaf(x, flog(me * rT), lgd));
  const bool zero = x < eps;
  const float a = fmaf(g.theta, d_log, -pi);
  const float spn = fmaxf(-pi, 0.f) + flog(1.f + ex2(-fabsf(pi) * LOG2E));  // softplus(-pi)
  const float oa = 1.f + ex2(-fabsf(a) * LOG2E);
  const float lsel = flog(zero ? oa : me * rT);
  const float ll = (zero ? fmaxf(a, 0.f) : a) - spn + (zero ? 1.f : x) * lsel + lgd;
  return x == eps ? 0.f : ll;  // x == eps: both reference masks are 0
}

// d(ll)/d(mu), d(ll)/d(theta), d(ll)/d(pi)
template <bool ZINB>
SCVI_FM void grad_mu(float x, float mu, float pi, const GeneK& g, float eps, float& dmu, float& dth, float& dpi) {
  const float me = mu + eps;
  const float rT = rcp(g.theta + mu + eps);
  const float d_log = -flog(fmaf(mu, g.rth, 1.f));
  const float da_dth = d_log + g.theta * (g.rth - rT);
  const float dgd = x == 0.f ? 0.f : digamma_pos_fast(x + g.theta) - g.dg_theta;
  if (!ZINB) {
    dmu = fmaf(x, rcp(me) - rT, -g.theta * rT);
    dth = da_dth - x * rT + dgd;
    dpi = 0.f;
    return;
  }
  const bool zero = x < eps;
  const float a = fmaf(g.theta, d_log, -pi);
  const float ep = ex2(-fabsf(pi) * LOG2E);
  const float rp = rcp(1.f + ep);
  const float ea = ex2(-fabsf(a) * LOG2E);
  const float r1 = rcp(zero ? 1.f + ea : me);
  const float sa = a >= 0.f ? r1 : ea * r1;  // sigmoid(a) when zero
  const float e_rp = ep * rp;
  dmu = zero ? -sa * g.theta * rT : fmaf(x, r1 - rT, -g.theta * rT);
  dth = zero ? sa * da_dth : da_dth - x * rT + dgd;
  dpi = zero ? (pi >= 0.f ? e_rp : rp) - sa : -(pi >= 0.f ? rp : e_rp);
  if (x == eps) dmu = dth = dpi = 0.f;
}

// One element of the backward: g1 = -mu * d(ll)/d(mu), dpi = d(ll)/d(pi), dth = d(ll)/d(theta), mu.
template <bool ZINB>
SCVI_FM void elem_bwd(float x, float s, float pi, float krow, const GeneK& g, float eps, float& g1, float& dpi,
                      float& dth, float& mu) {
  mu = ex2(fmaf(s, LOG2E, krow));
  const float T = g.theta + mu + eps;
  const float rT = rcp(T);
  const float d_log = -flog(fmaf(mu, g.rth, 1.f));
  const float da_dth = d_log + g.theta * (g.rth - rT);
  const float me = mu + eps;
  if (ZINB) {
    const float a = fmaf(g.theta, d_log, -pi);
    const float ep = ex2(-fabsf(pi) * LOG2E);
    const float rp = rcp(1.f + ep);
    const float ea = ex2(-fabsf(a) * LOG2E);
    const bool zero = x < eps;
    const float r1 = rcp(zero ? 1.f + ea : me);
    if (zero) {
      const float sa = a >= 0.f ? r1 : ea * r1;
      g1 = sa * g.theta * mu * rT;
      dpi = (pi >= 0.f ? ep * rp : rp) - sa;
      dth = sa * da_dth;
    } else if (x > eps) {
      g1 = g.theta * mu * rT * (1.f - x * r1);  // r1 - rT = theta / (me T)
      dpi = -(pi >= 0.f ? rp : ep * rp);
      dth = da_dth - x * rT + digamma_pos_fast(x + g.theta) - g.dg_theta;
    } else {
      g1 = dpi = dth = 0.f;
    }
  } else {
    g1 = g.theta * mu * rT;
    dpi = 0.f;
    dth = da_dth;
    if (x != 0.f) {
      g1 *= 1.f - x * rcp(me);
      dth += -x * rT + digamma_pos_fast(x + g.theta) - g.dg_theta;
    }
  }
}

// Table-driven, branch-free backward element (genes-owned dW kernel): dgd(x, theta) = digamma(x + theta) -
// digamma(theta) = sum_{k<x} 1/(theta + k) is read from a per-gene table of DGT entries (row stride DGT_LD, odd:
// conflict-free when a warp's lanes own consecutive genes); x >= DGT or non-integer x uses the series.
constexpr int DGT = 48, DGT_LD = 49;

template <bool ZINB>
SCVI_FM void elem_bwd_tab(float x, float s, float pi, float krow, const GeneK& g, float eps,
                          const float* __restrict__ tab /*[DGT] of this gene*/, float& g1, float& dpi, float& dth,
                          float& mu) {
  mu = ex2(fmaf(s, LOG2E, krow));
  const float me = mu + eps;
  const float rT = rcp(g.theta + mu + eps);
  const float d_log = -flog(fmaf(mu, g.rth, 1.f));
  const float da_dth = d_log + g.theta * (g.rth - rT);
  int xi;
  float dgd;
  if (count_index(x, DGT, xi)) dgd = tab[xi];
  else dgd = digamma_pos_fast(x + g.theta) - g.dg_theta;  // rare
  const float base = g.theta * mu * rT;
  const float nz_dth = da_dth - x * rT + dgd;
  if (ZINB) {
    const bool zero = x < eps;
    const float a = fmaf(g.theta, d_log, -pi);
    const float ep = ex2(-fabsf(pi) * LOG2E);
    const float rp = rcp(1.f + ep);
    const float ea = ex2(-fabsf(a) * LOG2E);
    const float r1 = rcp(zero ? 1.f + ea : me);
    const float sa = a >= 0.f ? r1 : ea * r1;  // sigmoid(a) when zero
    const float e_rp = ep * rp;
    g1 = base * (zero ? sa : fmaf(-x, r1, 1.f));
    dpi = zero ? (pi >= 0.f ? e_rp : rp) - sa : -(pi >= 0.f ? rp : e_rp);
    dth = zero ? sa * da_dth : nz_dth;
  } else {
    g1 = base * fmaf(-x, rcp(me), 1.f);
    dpi = 0.f;
    dth = nz_dth;  // x == 0: dgd = 0 and x * rT = 0
  }
}

}  // namespace fm
}  // namespace scvi
