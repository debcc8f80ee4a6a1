// BatchNorm1d(momentum 0.01, eps 1e-3) + ReLU + Dropout of FCLayers._build_layer
// (reference src/scvi/nn/_base_components.py:102-115), forward and full backward.
// Statistics are computed as per-tile (mean, M2) partials merged with Chan's parallel
// variance formula: deterministic, two-pass accuracy inside a tile, no atomics.
#include "common.cuh"

namespace scvi {

// ---- statistics ------------------------------------------------------------------------
__global__ void bn_partials_kernel(const float* __restrict__ y, int64_t ldy, int64_t B, int64_t N,
                                   float* __restrict__ part) {
  // 64 threads = 64 columns; one block per (row tile, 64-column group)
  const int64_t n = (int64_t)blockIdx.y * blockDim.x + threadIdx.x;
  const int64_t r0 = (int64_t)blockIdx.x * SCVI_BN_TILE;
  const int cnt = (int)min((int64_t)SCVI_BN_TILE, B - r0);
  if (n >= N) return;
  float s = 0.f;
  for (int r = 0; r < cnt; ++r) s += y[(r0 + r) * ldy + n];
  const float mean = s / (float)cnt;
  float m2 = 0.f;
  for (int r = 0; r < cnt; ++r) {
    const float d = y[(r0 + r) * ldy + n] - mean;
    m2 = fmaf(d, d, m2);
  }
  part[((int64_t)blockIdx.x * 2 + 0) * N + n] = mean;
  part[((int64_t)blockIdx.x * 2 + 1) * N + n] = m2;
}

// 32 column lanes x 32 tile groups per block: each thread merges its tiles serially (Chan), thread row 0
// then merges the 32 group results in fixed order.
__global__ void __launch_bounds__(1024)
    bn_finalize_kernel(const float* __restrict__ part, int64_t B, int64_t N, int training, float eps,
                       float momentum, float* __restrict__ run_mean, float* __restrict__ run_var,
                       float* __restrict__ mean_out, float* __restrict__ rstd_out) {
  __shared__ float s_cnt[32][33], s_mean[32][33], s_m2[32][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int64_t n = (int64_t)blockIdx.x * 32 + tx;
  if (!training) {
    if (ty == 0 && n < N) {
      mean_out[n] = run_mean[n];
      rstd_out[n] = 1.f / sqrtf(run_var[n] + eps);
    }
    return;
  }
  const int64_t T = (B + SCVI_BN_TILE - 1) / SCVI_BN_TILE;
  float cnt = 0.f, mean = 0.f, m2 = 0.f;
  if (n < N) {
    for (int64_t t = ty; t < T; t += 32) {
      const float cb = (float)min((int64_t)SCVI_BN_TILE, B - t * SCVI_BN_TILE);
      const float mb = part[(t * 2 + 0) * N + n];
      const float m2b = part[(t * 2 + 1) * N + n];
      const float tot = cnt + cb;
      const float delta = mb - mean;
      mean += delta * (cb / tot);
      m2 += m2b + delta * delta * (cnt * cb / tot);
      cnt = tot;
    }
  }
  s_cnt[ty][tx] = cnt;
  s_mean[ty][tx] = mean;
  s_m2[ty][tx] = m2;
  __syncthreads();
  if (ty != 0 || n >= N) return;
  cnt = 0.f; mean = 0.f; m2 = 0.f;
  for (int g = 0; g < 32; ++g) {
    const float cb = s_cnt[g][tx];
    if (cb == 0.f) continue;
    const float tot = cnt + cb;
    const float delta = s_mean[g][tx] - mean;
    mean += delta * (cb / tot);
    m2 += s_m2[g][tx] + delta * delta * (cnt * cb / tot);
    cnt = tot;
  }
  const float var_b = m2 / cnt;
  mean_out[n] = mean;
  rstd_out[n] = 1.f / sqrtf(var_b + eps);
  if (run_mean) run_mean[n] = (1.f - momentum) * run_mean[n] + momentum * mean;
  if (run_var) run_var[n] = (1.f - momentum) * run_var[n] + momentum * (m2 / fmaxf(cnt - 1.f, 1.f));
}

// ---- forward apply -----------------------------------------------------------------------
__global__ void bn_act_drop_fwd_kernel(const float* __restrict__ y, int64_t ldy, int64_t B, int64_t N,
                                       const float* __restrict__ mean, const float* __restrict__ rstd,
                                       const float* __restrict__ gamma, const float* __restrict__ beta, int relu,
                                       const uint8_t* __restrict__ mask, float drop_scale, float* __restrict__ a,
                                       int64_t lda) {
  const int64_t total = B * N;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total;
       e += (int64_t)gridDim.x * blockDim.x) {
    const int64_t b = e / N, n = e % N;
    float v = y[b * ldy + n];
    if (mean) {
      const float g = gamma ? gamma[n] : 1.f;
      v = (v - mean[n]) * (rstd[n] * g) + (beta ? beta[n] : 0.f);
    }
    if (relu) v = fmaxf(v, 0.f);
    if (mask) v = mask[e] ? v * drop_scale : 0.f;
    a[b * lda + n] = v;
  }
}

// ---- backward ------------------------------------------------------------------------------
// du = da * keep*scale * 1[u>0];  partial column sums of du and du*xhat per row tile
constexpr int BNR_RG = 4;  // row groups per 64-row tile (blockDim.y)
__global__ void __launch_bounds__(64 * BNR_RG)
    bn_bwd_reduce_kernel(const float* __restrict__ da, int64_t ldda, const float* __restrict__ y, int64_t ldy,
                         int64_t B, int64_t N, const float* __restrict__ mean, const float* __restrict__ rstd,
                         const float* __restrict__ gamma, const float* __restrict__ beta, int relu,
                         const uint8_t* __restrict__ mask, float drop_scale, float* __restrict__ part /*[T][2][N]*/) {
  // blockDim = (64 columns, BNR_RG row groups): group ty walks rows ty * 16 .. + 16 of the tile, eight rows per
  // memory round trip; the tile's partial is the sum of the groups' partials in group order (fixed).  Few-cell
  // minibatches launch a handful of these CTAs, so the per-thread serial chain is what sets the kernel's duration.
  __shared__ float sh[BNR_RG][2][64];
  constexpr int RPG = SCVI_BN_TILE / BNR_RG;
  const int64_t n = (int64_t)blockIdx.y * 64 + threadIdx.x;
  const int64_t r0 = (int64_t)blockIdx.x * SCVI_BN_TILE + (int64_t)threadIdx.y * RPG;
  const int cnt = (int)max((int64_t)0, min((int64_t)RPG, B - r0));
  const bool col_ok = n < N;
  float s1 = 0.f, s2 = 0.f;
  if (col_ok && cnt > 0) {
    const float mu = mean[n], rs = rstd[n];
    const float g = gamma ? gamma[n] : 1.f, bt = beta ? beta[n] : 0.f;
    for (int rg = 0; rg < cnt; rg += 8) {
      float yv[8], dv[8];
      uint8_t mv[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int64_t b = r0 + min(rg + i, cnt - 1);
        yv[i] = y[b * ldy + n];
        dv[i] = da[b * ldda + n];
        mv[i] = mask ? mask[b * N + n] : (uint8_t)1;
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) {  // straight-line: a row past the group contributes an exact zero
        const float xhat = (yv[i] - mu) * rs;
        const float u = xhat * g + bt;
        float d = dv[i];
        if (mask) d = mv[i] ? d * drop_scale : 0.f;
        if (relu && !(u > 0.f)) d = 0.f;
        if (rg + i >= cnt) d = 0.f;
        s1 += d;
        s2 = fmaf(d, xhat, s2);
      }
    }
  }
  sh[threadIdx.y][0][threadIdx.x] = s1;
  sh[threadIdx.y][1][threadIdx.x] = s2;
  __syncthreads();
  if (threadIdx.y == 0 && col_ok) {
    float t1 = 0.f, t2 = 0.f;
#pragma unroll
    for (int q = 0; q < BNR_RG; ++q) {
      t1 += sh[q][0][threadIdx.x];
      t2 += sh[q][1][threadIdx.x];
    }
    part[((int64_t)blockIdx.x * 2 + 0) * N + n] = t1;
    part[((int64_t)blockIdx.x * 2 + 1) * N + n] = t2;
  }
}

__global__ void bn_bwd_apply_kernel(const float* __restrict__ da, int64_t ldda, const float* __restrict__ y,
                                    int64_t ldy, int64_t B, int64_t N, const float* __restrict__ mean,
                                    const float* __restrict__ rstd, const float* __restrict__ gamma,
                                    const float* __restrict__ beta, int relu, const uint8_t* __restrict__ mask,
                                    float drop_scale, int training, const float* __restrict__ sums,
                                    float* __restrict__ dy, int64_t lddy) {
  const int64_t total = B * N;
  const float invB = 1.f / (float)B;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total;
       e += (int64_t)gridDim.x * blockDim.x) {
    const int64_t b = e / N, n = e % N;
    float d = da[b * ldda + n];
    if (mask) d = mask[e] ? d * drop_scale : 0.f;
    if (!mean) {
      if (relu && !(y[b * ldy + n] > 0.f)) d = 0.f;
      dy[b * lddy + n] = d;
      continue;
    }
    const float g = gamma ? gamma[n] : 1.f;
    const float xhat = (y[b * ldy + n] - mean[n]) * rstd[n];
    const float u = xhat * g + (beta ? beta[n] : 0.f);
    if (relu && !(u > 0.f)) d = 0.f;
    float r;
    if (training)
      r = rstd[n] * g * (d - sums[n] * invB - xhat * (sums[N + n] * invB));
    else
      r = rstd[n] * g * d;
    dy[b * lddy + n] = r;
  }
}

// ---------------------------------------------------------------------------------------
// Scratch of the allocation-free entry point scvi_bn_act_drop_bwd_static: lives in the library image
// (every GPU that loads the library holds these bytes until the process exits).  Layout = the `ws` of
// scvi_bn_act_drop_bwd: [ceil(B/64)][2][N] per-tile partial sums, then [2][N] column sums.
// In HOST code `g_bn_static_scratch` is only a stand-in for the symbol, not a device address: the
// address a kernel may use comes from cudaGetSymbolAddress, looked up once per device below.
// ---------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_bn_static_scratch[SCVI_BN_STATIC_SCRATCH_BYTES];

static int bn_static_scratch_ptr(void** out) {
  static std::atomic<void*> cache[64];  // per device ordinal; zero-initialised
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) {
    set_error("scvi_bn_act_drop_bwd_static: cudaGetDevice: %s", cudaGetErrorString(e));
    return SCVI_ECUDA;
  }
  if (dev < 0 || dev >= 64) {
    set_error("scvi_bn_act_drop_bwd_static: device ordinal %d out of range", dev);
    return SCVI_EUNSUPPORTED;
  }
  void* p = cache[dev].load(std::memory_order_acquire);
  if (!p) {  // first call on this device (which is current): resolve the symbol for it
    e = cudaGetSymbolAddress(&p, g_bn_static_scratch);
    if (e != cudaSuccess || !p) {
      cudaGetLastError();
      set_error("scvi_bn_act_drop_bwd_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
      return SCVI_ECUDA;
    }
    cache[dev].store(p, std::memory_order_release);  // two racing threads store the same address
  }
  *out = p;
  return SCVI_OK;
}

}  // namespace scvi

extern "C" {

int scvi_bn_partials(const float* y, int64_t ldy, int64_t B, int64_t N, float* bn_partials, void* stream) {
  using namespace scvi;
  SCVI_REQUIRE(y && bn_partials && B > 0 && N > 0 && ldy >= N, "scvi_bn_partials: bad argument");
  dim3 grid((unsigned)cdiv(B, SCVI_BN_TILE), (unsigned)cdiv(N, 64));
  bn_partials_kernel<<<grid, 64, 0, (cudaStream_t)stream>>>(y, ldy, B, N, bn_partials);
  SCVI_LAUNCH_CHECK("scvi_bn_partials");
  return SCVI_OK;
}

int scvi_bn_finalize(const float* bn_partials, int64_t B, int64_t N, int training, float eps, float momentum,
                     float* running_mean, float* running_var, float* mean, float* rstd, void* stream) {
  using namespace scvi;
  SCVI_REQUIRE(mean && rstd && B > 0 && N > 0, "scvi_bn_finalize: bad argument");
  SCVI_REQUIRE(training ? (bn_partials != nullptr) : (running_mean && running_var),
               "scvi_bn_finalize: missing partials (training) or running buffers (eval)");
  bn_finalize_kernel<<<(unsigned)cdiv(N, 32), 1024, 0, (cudaStream_t)stream>>>(
      bn_partials, B, N, training, eps, momentum, running_mean, running_var, mean, rstd);
  SCVI_LAUNCH_CHECK("scvi_bn_finalize");
  return SCVI_OK;
}

int scvi_bn_act_drop_fwd(const float* y, int64_t ldy, int64_t B, int64_t N, const float* mean, const float* rstd,
                         const float* gamma, const float* beta, int relu, const uint8_t* drop_mask,
                         float drop_scale, float* a, int64_t lda, void* stream) {
  using namespace scvi;
  SCVI_REQUIRE(y && a && B >= 0 && N > 0 && ldy >= N && lda >= N, "scvi_bn_act_drop_fwd: bad argument");
  SCVI_REQUIRE((mean == nullptr) == (rstd == nullptr), "scvi_bn_act_drop_fwd: mean and rstd go together");
  if (B == 0) return SCVI_OK;
  const int64_t total = B * N;
  const unsigned blocks = (unsigned)std::min<int64_t>(cdiv(total, 256), 148 * 8);
  bn_act_drop_fwd_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(y, ldy, B, N, mean, rstd, gamma, beta, relu,
                                                                   drop_mask, drop_scale, a, lda);
  SCVI_LAUNCH_CHECK("scvi_bn_act_drop_fwd");
  return SCVI_OK;
}

int64_t scvi_bn_bwd_ws_bytes(int64_t B, int64_t N) {
  return 4 * (scvi::cdiv(B, SCVI_BN_TILE) * 2 * N + 2 * N) + 256;
}

int scvi_bn_act_drop_bwd(const float* da, int64_t ldda, const float* y, int64_t ldy, int64_t B, int64_t N,
                         const float* mean, const float* rstd, const float* gamma, const float* beta, int relu,
                         const uint8_t* drop_mask, float drop_scale, int training, float* dy, int64_t lddy,
                         float* dgamma, float* dbeta, void* ws, void* stream) {
  using namespace scvi;
  SCVI_REQUIRE(da && y && dy && ws && B > 0 && N > 0, "scvi_bn_act_drop_bwd: bad argument");
  SCVI_REQUIRE((mean == nullptr) == (rstd == nullptr), "scvi_bn_act_drop_bwd: mean and rstd go together");
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t T = cdiv(B, SCVI_BN_TILE);
  float* part = (float*)ws;
  float* sums = part + T * 2 * N;
  if (mean) {
    dim3 grid((unsigned)T, (unsigned)cdiv(N, 64));
    bn_bwd_reduce_kernel<<<grid, dim3(64, BNR_RG), 0, st>>>(da, ldda, y, ldy, B, N, mean, rstd, gamma, beta, relu, drop_mask,
                                              drop_scale, part);
    SCVI_LAUNCH_CHECK("scvi_bn_act_drop_bwd(reduce)");
    // sums[0..N) = sum_t du, sums[N..2N) = sum_t du * xhat; the same rows are dbeta / dgamma
    ColsumSegs segs{};
    segs.n = 3;
    segs.dst[0] = sums; segs.begin[0] = 0; segs.end[0] = 2 * N;
    segs.dst[1] = dbeta; segs.begin[1] = 0; segs.end[1] = N;
    segs.dst[2] = dgamma; segs.begin[2] = N; segs.end[2] = 2 * N;
    int rc = launch_colsum(part, T, 2 * N, 2 * N, segs, st, "scvi_bn_act_drop_bwd(sums)");
    if (rc) return rc;
  }
  const int64_t total = B * N;
  const unsigned blocks = (unsigned)std::min<int64_t>(cdiv(total, 256), 148 * 8);
  bn_bwd_apply_kernel<<<blocks, 256, 0, st>>>(da, ldda, y, ldy, B, N, mean, rstd, gamma, beta, relu, drop_mask,
                                              drop_scale, training, sums, dy, lddy);
  SCVI_LAUNCH_CHECK("scvi_bn_act_drop_bwd(apply)");
  return SCVI_OK;
}

int64_t scvi_bn_static_scratch_bytes(void) { return (int64_t)SCVI_BN_STATIC_SCRATCH_BYTES; }

int scvi_bn_act_drop_bwd_static(const float* da, int64_t ldda, const float* y, int64_t ldy, int64_t B, int64_t N,
                                const float* mean, const float* rstd, const float* gamma, const float* beta,
                                int relu, const uint8_t* drop_mask, float drop_scale, int training, float* dy,
                                int64_t lddy, float* dgamma, float* dbeta, void* stream) {
  using namespace scvi;
  SCVI_REQUIRE(B > 0 && N > 0, "scvi_bn_act_drop_bwd_static: bad shape");
  const int64_t need = scvi_bn_bwd_ws_bytes(B, N);
  if (need > (int64_t)SCVI_BN_STATIC_SCRATCH_BYTES) {
    set_error("scvi_bn_act_drop_bwd_static: B=%lld x N=%lld needs %lld bytes of scratch, the built-in scratch holds "
              "%lld (use scvi_bn_act_drop_bwd with a caller-owned workspace)",
              (long long)B, (long long)N, (long long)need, (long long)SCVI_BN_STATIC_SCRATCH_BYTES);
    return SCVI_EINVAL;
  }
  void* scratch = nullptr;
  if (int rc = bn_static_scratch_ptr(&scratch)) return rc;
  // every partial and both column sums are overwritten before they are read: nothing has to start at zero
  return scvi_bn_act_drop_bwd(da, ldda, y, ldy, B, N, mean, rstd, gamma, beta, relu, drop_mask, drop_scale, training,
                              dy, lddy, dgamma, dbeta, scratch, stream);
}
}
