// Host build of scvi_tools_b200/csrc/fastmath.cuh (exp2f / log2f stand in for the SFU ops) so the
// per-element formulas are checked against float64 without a GPU.  Test infrastructure only.
#include <cmath>
#include <vector>
#define SCVI_FASTMATH_HOST 1
#include "../../scvi_tools_b200/csrc/fastmath.cuh"

using namespace scvi::fm;

static GeneK make_gene(float theta, float eps) {
  GeneK g;
  g.theta = theta;
  g.log_theta = (float)std::log((double)(theta + eps));
  g.lg_theta = (float)std::lgamma((double)theta);
  g.rth = 1.0f / (theta + eps);
  // digamma by finite recurrence + asymptotic series in double
  double z = theta, acc = 0.0;
  while (z < 10.0) { acc -= 1.0 / z; z += 1.0; }
  const double r = 1.0 / z, r2 = r * r;
  acc += std::log(z) - 0.5 * r - r2 * (1.0 / 12 - r2 * (1.0 / 120 - r2 * (1.0 / 252 - r2 * (1.0 / 240 - r2 / 132))));
  g.dg_theta = (float)acc;
  return g;
}

extern "C" {
void fm_ex2_fma(int n, const float* x, float* y) {
  for (int i = 0; i < n; ++i) y[i] = scvi::fm::ex2_fma(x[i]);
}

// out[n][4] = ll, g1, dpi, mu
void fm_elem_fwd(int zinb, int n, const float* x, const float* s, const float* pi, const float* krow,
                 const float* theta, float eps, float* out) {
  std::vector<float> tab(LGX1_TABLE);
  for (int i = 0; i < LGX1_TABLE; ++i) tab[i] = (float)std::lgamma((double)i + 1.0);
  for (int i = 0; i < n; ++i) {
    const GeneK g = make_gene(theta[i], eps);
    float ll, g1, dpi, mu;
    if (zinb) elem_fwd<true>(x[i], s[i], pi[i], krow[i], g, eps, tab.data(), ll, g1, dpi, mu);
    else elem_fwd<false>(x[i], s[i], pi[i], krow[i], g, eps, tab.data(), ll, g1, dpi, mu);
    out[4 * i] = ll; out[4 * i + 1] = g1; out[4 * i + 2] = dpi; out[4 * i + 3] = mu;
  }
}
// table-driven branch-free variant (the per-gene table is built here the way the pack kernel builds it)
void fm_elem_fwd_tab(int zinb, int n, const float* x, const float* s, const float* pi, const float* krow,
                     const float* theta, float eps, float* out) {
  std::vector<float> lgx1(LGX1_TABLE), tab(LGT);
  for (int i = 0; i < LGX1_TABLE; ++i) lgx1[i] = (float)std::lgamma((double)i + 1.0);
  for (int i = 0; i < n; ++i) {
    const GeneK g = make_gene(theta[i], eps);
    for (int k = 0; k < LGT; ++k)
      tab[k] = k == 0 ? 0.f : (float)(std::lgamma((double)k + (double)theta[i]) - std::lgamma((double)theta[i]) - std::lgamma(k + 1.0));
    float ll, g1, dpi, mu;
    if (zinb) elem_fwd_tab<true>(x[i], s[i], pi[i], krow[i], g, eps, tab.data(), lgx1.data(), ll, g1, dpi, mu);
    else elem_fwd_tab<false>(x[i], s[i], pi[i], krow[i], g, eps, tab.data(), lgx1.data(), ll, g1, dpi, mu);
    out[4 * i] = ll; out[4 * i + 1] = g1; out[4 * i + 2] = dpi; out[4 * i + 3] = mu;
  }
}
// out[n][4] = g1, dpi, dth, mu
void fm_elem_bwd(int zinb, int n, const float* x, const float* s, const float* pi, const float* krow,
                 const float* theta, float eps, float* out) {
  for (int i = 0; i < n; ++i) {
    const GeneK g = make_gene(theta[i], eps);
    float g1, dpi, dth, mu;
    if (zinb) elem_bwd<true>(x[i], s[i], pi[i], krow[i], g, eps, g1, dpi, dth, mu);
    else elem_bwd<false>(x[i], s[i], pi[i], krow[i], g, eps, g1, dpi, dth, mu);
    out[4 * i] = g1; out[4 * i + 1] = dpi; out[4 * i + 2] = dth; out[4 * i + 3] = mu;
  }
}
// table-driven branch-free backward element; the dgd table is built as the pack kernel builds it
void fm_elem_bwd_tab(int zinb, int n, const float* x, const float* s, const float* pi, const float* krow,
                     const float* theta, float eps, float* out) {
  std::vector<float> tab(DGT_LD);
  for (int i = 0; i < n; ++i) {
    const GeneK g = make_gene(theta[i], eps);
    double acc = 0.0;
    for (int k = 0; k < DGT_LD; ++k) {
      tab[k] = (float)acc;
      if (k < DGT) acc += 1.0 / ((double)theta[i] + k);
    }
    float g1, dpi, dth, mu;
    if (zinb) elem_bwd_tab<true>(x[i], s[i], pi[i], krow[i], g, eps, tab.data(), g1, dpi, dth, mu);
    else elem_bwd_tab<false>(x[i], s[i], pi[i], krow[i], g, eps, tab.data(), g1, dpi, dth, mu);
    out[4 * i] = g1; out[4 * i + 1] = dpi; out[4 * i + 2] = dth; out[4 * i + 3] = mu;
  }
}
// lean straight-line variants (integer x below the table length): out[n][4] = ll, g1, dpi, mu / g1, dpi, dth, mu
void fm_elem_fwd_lean(int zinb, int n, const float* x, const float* s, const float* pi, const float* krow,
                      const float* theta, float eps, float* out) {
  for (int i = 0; i < n; ++i) {
    const GeneK g = make_gene(theta[i], eps);
    const int k = (int)x[i];
    const float lgd = k == 0 ? 0.f : (float)(std::lgamma((double)k + (double)theta[i]) - std::lgamma((double)theta[i]) - std::lgamma(k + 1.0));
    float ll, g1, dpi, mu;
    if (zinb) elem_fwd_lean<true>(x[i], lgd, s[i], pi[i], krow[i], g, eps, ll, g1, dpi, mu);
    else elem_fwd_lean<false>(x[i], lgd, s[i], pi[i], krow[i], g, eps, ll, g1, dpi, mu);
    out[4 * i] = ll; out[4 * i + 1] = g1; out[4 * i + 2] = dpi; out[4 * i + 3] = mu;
  }
}
void fm_elem_bwd_lean(int zinb, int n, const float* x, const float* s, const float* pi, const float* krow,
                      const float* theta, float eps, float* out) {
  for (int i = 0; i < n; ++i) {
    const GeneK g = make_gene(theta[i], eps);
    const int k = (int)x[i];
    double acc = 0.0;
    for (int j = 0; j < k; ++j) acc += 1.0 / ((double)theta[i] + j);
    float g1, dpi, dth, mu;
    if (zinb) elem_bwd_lean<true>(x[i], (float)acc, s[i], pi[i], krow[i], g, eps, g1, dpi, dth, mu);
    else elem_bwd_lean<false>(x[i], (float)acc, s[i], pi[i], krow[i], g, eps, g1, dpi, dth, mu);
    out[4 * i] = g1; out[4 * i + 1] = dpi; out[4 * i + 2] = dth; out[4 * i + 3] = mu;
  }
}
void fm_lgamma(int n, const float* z, float* out) { for (int i = 0; i < n; ++i) out[i] = lgamma_pos(z[i]); }
void fm_digamma(int n, const float* z, float* out) { for (int i = 0; i < n; ++i) out[i] = digamma_pos_fast(z[i]); }
}
