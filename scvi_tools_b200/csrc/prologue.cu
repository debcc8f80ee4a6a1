// Step prologue: everything the minibatch step needs that does not depend on the counts, in one launch.
//   * dropout keep-mask of the encoder hidden layer   (nn.Dropout(p) in FCLayers,
//     reference src/scvi/nn/_base_components.py:95-104 -- Bernoulli(1-p) per element)
//   * standard-normal noise of the reparameterised draw (Normal(q_m, sqrt(q_v)).rsample(),
//     src/scvi/nn/_base_components.py:286-289)
//   * one-hot batch covariate as a weight-column index  (cov_cols[b] = batch_cols[idx[b]],
//     the one_hot + cat of FCLayers.forward, src/scvi/nn/_base_components.py:139-169)
//   * the backward seeds of the ELBO mean              (src/scvi/module/_vae.py:573-580)
//   * BatchNorm num_batches_tracked counters
// Random numbers: Philox4x32-10 keyed by `seed`, counter = (element block, step, stream id, domain), with
// the step read from device memory so a captured CUDA graph draws fresh numbers at every replay.
// The streams are this library's own (they do not reproduce torch's generator): parity tests feed explicit
// masks and noise, statistical tests check the moments.
#include <algorithm>

#include "common.cuh"

namespace scvi {

struct Philox {
  uint32_t k0, k1;
  __device__ __forceinline__ uint4 operator()(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3) const {
    uint32_t a = k0, b = k1;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
      const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
      const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
      c0 = hi1 ^ c1 ^ a;
      c1 = lo1;
      c2 = hi0 ^ c3 ^ b;
      c3 = lo0;
      a += 0x9E3779B9u;
      b += 0xBB67AE85u;
    }
    return make_uint4(c0, c1, c2, c3);
  }
};

__device__ __forceinline__ float u01_open(uint32_t u) { return ((float)(u >> 8) + 0.5f) * (1.f / 16777216.f); }  // (0,1)

__global__ void __launch_bounds__(256)
    step_prologue_kernel(uint64_t seed, const int64_t* __restrict__ step_counter, uint32_t stream_id, int64_t B,
                         int64_t n_mask /*B*H*/, int64_t n_noise /*B*L*/, float p_drop, uint8_t* __restrict__ mask,
                         float* __restrict__ noise, const int64_t* __restrict__ idx,
                         const int32_t* __restrict__ batch_cols, int32_t* __restrict__ cov_cols, float klw_host,
                         const float* __restrict__ klw_dev, float scale, float* __restrict__ g_reconst,
                         float* __restrict__ g_kl, int64_t* __restrict__ nbt1, int64_t* __restrict__ nbt2) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t step = step_counter ? *step_counter : 0;
  Philox ph{(uint32_t)seed, (uint32_t)(seed >> 32)};
  // counter words: (element block lo, element block hi | domain << 28, step lo, step hi16 | stream id << 16)
  const uint32_t c2 = (uint32_t)step, c3 = ((uint32_t)(step >> 32) & 0xFFFFu) | (stream_id << 16);
  // ---- keep-mask: 16 elements per thread (4 Philox calls), keep <=> u32 >= p * 2^32
  if (mask) {
    const int64_t e0 = t * 16;
    if (e0 < n_mask) {
      const uint32_t thr = (uint32_t)fminf(p_drop * 4294967296.f, 4294967040.f);
      uint32_t w[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int64_t i = 4 * t + q;
        const uint4 r = ph((uint32_t)i, (uint32_t)(i >> 32), c2, c3);  // domain 0
        w[q] = (r.x >= thr ? 1u : 0u) | (r.y >= thr ? 0x100u : 0u) | (r.z >= thr ? 0x10000u : 0u) | (r.w >= thr ? 0x1000000u : 0u);
      }
      if (e0 + 16 <= n_mask) {
        *reinterpret_cast<uint4*>(mask + e0) = make_uint4(w[0], w[1], w[2], w[3]);
      } else {
        for (int i = 0; e0 + i < n_mask; ++i) mask[e0 + i] = (uint8_t)((w[i >> 2] >> (8 * (i & 3))) & 1u);
      }
    }
  }
  // ---- N(0,1) noise: 4 values per thread (Box-Muller on one Philox call)
  if (noise) {
    const int64_t e0 = t * 4;
    if (e0 < n_noise) {
      const uint4 r = ph((uint32_t)t, (uint32_t)(t >> 32) | 0x10000000u, c2, c3);  // domain 1
      const float r0 = sqrtf(-2.f * logf(u01_open(r.x))), r1 = sqrtf(-2.f * logf(u01_open(r.z)));
      float s0, c0, s1, c1;
      sincospif(2.f * u01_open(r.y), &s0, &c0);
      sincospif(2.f * u01_open(r.w), &s1, &c1);
      const float v[4] = {r0 * c0, r0 * s0, r1 * c1, r1 * s1};
      for (int i = 0; i < 4 && e0 + i < n_noise; ++i) noise[e0 + i] = v[i];
    }
  }
  if (t < B) {
    if (cov_cols) cov_cols[t] = batch_cols[idx ? idx[t] : t];
    const float klw = klw_dev ? *klw_dev : klw_host;
    const float s = scale / (float)B;
    if (g_reconst) g_reconst[t] = s;
    if (g_kl) g_kl[t] = s * klw;
  }
  if (t == 0) {
    if (nbt1) *nbt1 += 1;
    if (nbt2) *nbt2 += 1;
  }
}

}  // namespace scvi

extern "C" int scvi_step_prologue(uint64_t seed, const int64_t* step_counter, int64_t stream_id, int64_t B, int64_t H,
                                  int64_t L, float p_drop, uint8_t* mask, float* noise, const int64_t* idx,
                                  const int32_t* batch_cols, int32_t* cov_cols, float kl_weight,
                                  const float* kl_weight_dev, float scale, float* g_reconst, float* g_kl,
                                  int64_t* nbt1, int64_t* nbt2, void* stream) {
  using namespace scvi;
  SCVI_REQUIRE(B > 0 && H >= 0 && L >= 0, "scvi_step_prologue: bad shape");
  SCVI_REQUIRE(p_drop >= 0.f && p_drop < 1.f, "scvi_step_prologue: p_drop must be in [0, 1)");
  SCVI_REQUIRE(!cov_cols || batch_cols, "scvi_step_prologue: cov_cols needs batch_cols");
  SCVI_REQUIRE(!mask || ((uintptr_t)mask % 16 == 0), "scvi_step_prologue: mask must be 16-byte aligned");
  const int64_t n_mask = mask ? B * H : 0, n_noise = noise ? B * L : 0;
  const int64_t threads = std::max<int64_t>(std::max<int64_t>(cdiv(n_mask, 16), cdiv(n_noise, 4)), B);
  step_prologue_kernel<<<(unsigned)cdiv(threads, 256), 256, 0, (cudaStream_t)stream>>>(
      seed, step_counter, (uint32_t)stream_id, B, n_mask, n_noise, p_drop, mask, noise, idx, batch_cols, cov_cols,
      kl_weight, kl_weight_dev, scale, g_reconst, g_kl, nbt1, nbt2);
  SCVI_LAUNCH_CHECK("scvi_step_prologue");
  return SCVI_OK;
}
