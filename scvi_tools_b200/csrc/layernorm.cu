// LayerNorm (no affine) + ReLU + Dropout of FCLayers._build_layer when `use_layer_norm` is set, and the
// gradient of the batch embedding table.
//
// Replaces nn.LayerNorm(n_out, elementwise_affine=False) -> nn.ReLU -> nn.Dropout
// (reference src/scvi/nn/_base_components.py:102-115) and the backward of scvi.nn.Embedding
// (reference src/scvi/nn/_embedding.py:29-98; forward = a row gather of the table, scvi_gather_rows).
//
// LayerNorm normalises every ROW: one warp per row, two passes over the row in registers / L1 (mean, then the
// centred sum of squares: the same two-pass arithmetic torch uses, eps inside the square root), statistics
// combined with warp shuffles.  HBM-bound: 8 bytes per element forward, 12 backward.
#include "common.cuh"

namespace scvi {

constexpr int LN_WARPS = 8;  // rows per block

__global__ void __launch_bounds__(32 * LN_WARPS)
    ln_act_drop_fwd_kernel(const float* __restrict__ y, int64_t ldy, int64_t B, int64_t N, float eps, int relu,
                           const uint8_t* __restrict__ mask, float drop_scale, float* __restrict__ a, int64_t lda,
                           float* __restrict__ stats /*[B][2] = (mean, rstd), nullable*/) {
  const int lane = threadIdx.x & 31;
  const int64_t b = (int64_t)blockIdx.x * LN_WARPS + (threadIdx.x >> 5);
  if (b >= B) return;
  const float* row = y + b * ldy;
  float s = 0.f;
  for (int64_t n = lane; n < N; n += 32) s += row[n];
  const float mean = warp_sum(s) / (float)N;
  float q = 0.f;
  for (int64_t n = lane; n < N; n += 32) {
    const float d = row[n] - mean;
    q = fmaf(d, d, q);
  }
  const float rstd = 1.f / sqrtf(warp_sum(q) / (float)N + eps);
  if (stats && lane == 0) {
    stats[2 * b] = mean;
    stats[2 * b + 1] = rstd;
  }
  for (int64_t n = lane; n < N; n += 32) {
    float v = (row[n] - mean) * rstd;
    if (relu) v = fmaxf(v, 0.f);
    if (mask) v = mask[b * N + n] ? v * drop_scale : 0.f;
    a[b * lda + n] = v;
  }
}

// dy = rstd * (g - mean_n(g) - xhat * mean_n(g * xhat)),  g = da through dropout and ReLU
__global__ void __launch_bounds__(32 * LN_WARPS)
    ln_act_drop_bwd_kernel(const float* __restrict__ da, int64_t ldda, const float* __restrict__ y, int64_t ldy,
                           int64_t B, int64_t N, const float* __restrict__ stats, int relu,
                           const uint8_t* __restrict__ mask, float drop_scale, float* __restrict__ dy, int64_t lddy) {
  const int lane = threadIdx.x & 31;
  const int64_t b = (int64_t)blockIdx.x * LN_WARPS + (threadIdx.x >> 5);
  if (b >= B) return;
  const float mean = stats[2 * b], rstd = stats[2 * b + 1];
  const float* row = y + b * ldy;
  const float* grow = da + b * ldda;
  float s1 = 0.f, s2 = 0.f;
  for (int64_t n = lane; n < N; n += 32) {
    const float xhat = (row[n] - mean) * rstd;
    float g = grow[n];
    if (mask) g = mask[b * N + n] ? g * drop_scale : 0.f;
    if (relu && !(xhat > 0.f)) g = 0.f;
    s1 += g;
    s2 = fmaf(g, xhat, s2);
  }
  const float m1 = warp_sum(s1) / (float)N, m2 = warp_sum(s2) / (float)N;
  for (int64_t n = lane; n < N; n += 32) {
    const float xhat = (row[n] - mean) * rstd;
    float g = grow[n];
    if (mask) g = mask[b * N + n] ? g * drop_scale : 0.f;
    if (relu && !(xhat > 0.f)) g = 0.f;
    dy[b * lddy + n] = rstd * (g - m1 - xhat * m2);
  }
}

// dW[e, :] = sum_{b : idx[b] == e} g[b, :], rows visited in order (deterministic, no atomics).
// grid = (n_emb, ceil(D / 32)); block = 32 columns x 8 row groups; group r takes rows r, r + 8, ...
__global__ void __launch_bounds__(256)
    embedding_bwd_kernel(const float* __restrict__ g, int64_t ldg, const int64_t* __restrict__ idx, int64_t B, int64_t D,
                         float* __restrict__ dW) {
  __shared__ float s_part[8][33];
  const int64_t e = blockIdx.x;
  const int64_t d = (int64_t)blockIdx.y * 32 + (threadIdx.x & 31);
  const int rg = threadIdx.x >> 5;
  float acc = 0.f;
  if (d < D)
    for (int64_t b = rg; b < B; b += 8)
      if (idx[b] == e) acc += g[b * ldg + d];
  s_part[rg][threadIdx.x & 31] = acc;
  __syncthreads();
  if (rg == 0 && d < D) {
    float tot = 0.f;
#pragma unroll
    for (int r = 0; r < 8; ++r) tot += s_part[r][threadIdx.x & 31];
    dW[e * D + d] = tot;
  }
}

}  // namespace scvi

extern "C" {

int scvi_ln_act_drop_fwd(const float* y, int64_t ldy, int64_t B, int64_t N, float eps, int relu,
                         const uint8_t* drop_mask, float drop_scale, float* a, int64_t lda, float* stats,
                         void* stream) {
  using namespace scvi;
  SCVI_REQUIRE(y && a && B >= 0 && N > 0 && ldy >= N && lda >= N, "scvi_ln_act_drop_fwd: bad argument");
  if (B == 0) return SCVI_OK;
  ln_act_drop_fwd_kernel<<<(unsigned)cdiv(B, LN_WARPS), 32 * LN_WARPS, 0, (cudaStream_t)stream>>>(
      y, ldy, B, N, eps, relu, drop_mask, drop_scale, a, lda, stats);
  SCVI_LAUNCH_CHECK("scvi_ln_act_drop_fwd");
  return SCVI_OK;
}

int scvi_ln_act_drop_bwd(const float* da, int64_t ldda, const float* y, int64_t ldy, int64_t B, int64_t N,
                         const float* stats, int relu, const uint8_t* drop_mask, float drop_scale, float* dy,
                         int64_t lddy, void* stream) {
  using namespace scvi;
  SCVI_REQUIRE(da && y && stats && dy && B >= 0 && N > 0 && ldda >= N && ldy >= N && lddy >= N,
               "scvi_ln_act_drop_bwd: bad argument");
  if (B == 0) return SCVI_OK;
  ln_act_drop_bwd_kernel<<<(unsigned)cdiv(B, LN_WARPS), 32 * LN_WARPS, 0, (cudaStream_t)stream>>>(
      da, ldda, y, ldy, B, N, stats, relu, drop_mask, drop_scale, dy, lddy);
  SCVI_LAUNCH_CHECK("scvi_ln_act_drop_bwd");
  return SCVI_OK;
}

int scvi_embedding_bwd(const float* g, int64_t ldg, const int64_t* idx, int64_t B, int64_t D, int64_t n_emb,
                       float* dW, void* stream) {
  using namespace scvi;
  SCVI_REQUIRE(g && idx && dW && B >= 0 && D > 0 && n_emb > 0 && ldg >= D, "scvi_embedding_bwd: bad argument");
  SCVI_REQUIRE(n_emb <= 65535, "scvi_embedding_bwd: at most 65535 embedding rows");
  dim3 grid((unsigned)n_emb, (unsigned)cdiv(D, 32));
  embedding_bwd_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(g, ldg, idx, B, D, dW);
  SCVI_LAUNCH_CHECK("scvi_embedding_bwd");
  return SCVI_OK;
}
}
