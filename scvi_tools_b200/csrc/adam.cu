// Multi-tensor Adam over one flat fp32 arena + ELBO reduction / seeding.
// Replaces torch.optim.Adam(lr=1e-3, eps=0.01, weight_decay=1e-6) as configured by
// TrainingPlan.configure_optimizers (reference src/scvi/train/_trainingplans.py:475-482,497-519)
// -- coupled L2, bias correction, op order of torch's single-tensor path -- and the
// mean over the minibatch in VAE.loss (src/scvi/module/_vae.py:573-580).
// HBM-bound: 28 bytes per parameter (read p, g, m, v; write p, m, v).
#include "common.cuh"

namespace scvi {

__device__ __forceinline__ void adam_elem(float& p, float g, float& m, float& v, float grad_scale, float wd,
                                          float beta1, float beta2, float step_size, float bc2_sqrt, float eps) {
  g = g * grad_scale;
  g = fmaf(wd, p, g);                          // grad.add(param, alpha=weight_decay)
  m = m + (g - m) * (1.f - beta1);             // exp_avg.lerp_(grad, 1 - beta1)
  v = v * beta2 + (1.f - beta2) * g * g;       // exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value=1-beta2)
  const float denom = sqrtf(v) / bc2_sqrt + eps;
  p = p - step_size * (m / denom);             // param.addcdiv_(exp_avg, denom, value=-step_size)
}

__global__ void adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                            float* __restrict__ v, int64_t n, int64_t* step_counter, int64_t step_host, float lr,
                            float beta1, float beta2, float eps, float wd, float grad_scale) {
  const int64_t step = step_counter ? (*reinterpret_cast<volatile int64_t*>(step_counter) + 1) : step_host;
  const double bc1 = 1.0 - pow((double)beta1, (double)step);
  const double bc2 = 1.0 - pow((double)beta2, (double)step);
  const float step_size = (float)((double)lr / bc1);
  const float bc2_sqrt = (float)sqrt(bc2);
  const int64_t n4 = n >> 2;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    float4 pp = reinterpret_cast<float4*>(p)[i];
    const float4 gg = reinterpret_cast<const float4*>(g)[i];
    float4 mm = reinterpret_cast<float4*>(m)[i];
    float4 vv = reinterpret_cast<float4*>(v)[i];
    adam_elem(pp.x, gg.x, mm.x, vv.x, grad_scale, wd, beta1, beta2, step_size, bc2_sqrt, eps);
    adam_elem(pp.y, gg.y, mm.y, vv.y, grad_scale, wd, beta1, beta2, step_size, bc2_sqrt, eps);
    adam_elem(pp.z, gg.z, mm.z, vv.z, grad_scale, wd, beta1, beta2, step_size, bc2_sqrt, eps);
    adam_elem(pp.w, gg.w, mm.w, vv.w, grad_scale, wd, beta1, beta2, step_size, bc2_sqrt, eps);
    reinterpret_cast<float4*>(p)[i] = pp;
    reinterpret_cast<float4*>(m)[i] = mm;
    reinterpret_cast<float4*>(v)[i] = vv;
  }
  for (int64_t i = 4 * n4 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    adam_elem(p[i], g[i], m[i], v[i], grad_scale, wd, beta1, beta2, step_size, bc2_sqrt, eps);
  // device-side step counter: every block has read it above, so the last block to finish may bump it
  // (step_counter[1] is the arrival ticket)
  if (step_counter) {
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence();
      unsigned long long* ticket = reinterpret_cast<unsigned long long*>(step_counter + 1);
      if (atomicAdd(ticket, 1ull) == (unsigned long long)gridDim.x - 1) {
        *ticket = 0ull;
        step_counter[0] = step;
      }
    }
  }
}

// out[0] = mean(reconst + klw * kl), out[1] = sum reconst, out[2] = sum kl  (single block, fixed order)
__global__ void elbo_reduce_kernel(const float* __restrict__ reconst, const float* __restrict__ kl, int64_t B,
                                   float klw_host, const float* __restrict__ klw_dev, float* __restrict__ out) {
  __shared__ float sh[3][32];
  const float klw = klw_dev ? *klw_dev : klw_host;
  float a = 0.f, r = 0.f, k = 0.f;
  for (int64_t i = threadIdx.x; i < B; i += blockDim.x) {
    const float rv = reconst[i], kv = kl[i];
    a += rv + klw * kv;
    r += rv;
    k += kv;
  }
  a = warp_sum(a);
  r = warp_sum(r);
  k = warp_sum(k);
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (lane == 0) {
    sh[0][w] = a;
    sh[1][w] = r;
    sh[2][w] = k;
  }
  __syncthreads();
  if (w == 0) {
    const int nw = blockDim.x >> 5;
    a = lane < nw ? sh[0][lane] : 0.f;
    r = lane < nw ? sh[1][lane] : 0.f;
    k = lane < nw ? sh[2][lane] : 0.f;
    a = warp_sum(a);
    r = warp_sum(r);
    k = warp_sum(k);
    if (lane == 0) {
      out[0] = a / (float)B;
      out[1] = r;
      out[2] = k;
    }
  }
}

__global__ void elbo_seed_kernel(int64_t B, float klw_host, const float* __restrict__ klw_dev, float scale,
                                 float* __restrict__ g_reconst, float* __restrict__ g_kl) {
  const float klw = klw_dev ? *klw_dev : klw_host;
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < B) {
    const float s = scale / (float)B;
    g_reconst[i] = s;
    g_kl[i] = s * klw;
  }
}

// Gradient tensors scattered over the allocator -> the flat gradient arena, one launch for up to GS_MAX tensors
// (the table travels by value in the launch parameters: captured into a CUDA graph with the step, no device-side
// pointer table to keep alive).  A NULL source zero-fills its slot (a parameter that received no gradient).
constexpr int GS_MAX = 64;
struct GatherSegs {
  const float* src[GS_MAX];
  int64_t off[GS_MAX];
  int64_t n[GS_MAX];
};
__global__ void gather_segments_kernel(const GatherSegs s, float* __restrict__ dst) {
  const int seg = blockIdx.y;
  const float* __restrict__ src = s.src[seg];
  float* __restrict__ d = dst + s.off[seg];
  const int64_t n = s.n[seg];
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t t0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (src == nullptr) {
    for (int64_t i = t0; i < n; i += stride) d[i] = 0.f;
    return;
  }
  if ((((uintptr_t)src | (uintptr_t)d) & 15) == 0) {
    const int64_t n4 = n >> 2;
    for (int64_t i = t0; i < n4; i += stride) reinterpret_cast<float4*>(d)[i] = reinterpret_cast<const float4*>(src)[i];
    for (int64_t i = 4 * n4 + t0; i < n; i += stride) d[i] = src[i];
  } else {
    for (int64_t i = t0; i < n; i += stride) d[i] = src[i];
  }
}

static int adam_launch(float* p, const float* g, float* m, float* v, int64_t n, int64_t* counter, int64_t step,
                       float lr, float beta1, float beta2, float eps, float wd, float grad_scale, void* stream,
                       const char* name) {
  SCVI_REQUIRE(p && g && m && v && n >= 0, "%s: bad argument", name);
  SCVI_REQUIRE(((uintptr_t)p % 16 == 0) && ((uintptr_t)g % 16 == 0) && ((uintptr_t)m % 16 == 0) &&
                   ((uintptr_t)v % 16 == 0),
               "%s: arenas must be 16-byte aligned", name);
  SCVI_REQUIRE(counter || step >= 1, "%s: step must be >= 1", name);
  if (n == 0) return SCVI_OK;
  const unsigned blocks = (unsigned)std::min<int64_t>(cdiv(cdiv(n, 4), 256), 148 * 8);
  adam_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(p, g, m, v, n, counter, step, lr, beta1, beta2, eps, wd, grad_scale);
  SCVI_LAUNCH_CHECK(name);
  return SCVI_OK;
}

}  // namespace scvi

extern "C" {

int scvi_adam_step(float* p, const float* g, float* m, float* v, int64_t n, int64_t step, float lr, float beta1,
                   float beta2, float eps, float weight_decay, float grad_scale, void* stream) {
  return scvi::adam_launch(p, g, m, v, n, nullptr, step, lr, beta1, beta2, eps, weight_decay, grad_scale, stream, "scvi_adam_step");
}

int scvi_gather_segments(const void* const* src, const int64_t* offs, const int64_t* ns, int count, float* dst,
                         void* stream) {
  using namespace scvi;
  SCVI_REQUIRE(count >= 0 && (count == 0 || (src && offs && ns && dst)), "scvi_gather_segments: bad argument");
  for (int c0 = 0; c0 < count; c0 += GS_MAX) {
    const int c = std::min(GS_MAX, count - c0);
    GatherSegs s;
    int64_t nmax = 0;
    for (int i = 0; i < c; ++i) {
      SCVI_REQUIRE(offs[c0 + i] >= 0 && ns[c0 + i] >= 0, "scvi_gather_segments: negative offset or length");
      s.src[i] = (const float*)src[c0 + i];
      s.off[i] = offs[c0 + i];
      s.n[i] = ns[c0 + i];
      nmax = std::max(nmax, ns[c0 + i]);
    }
    if (nmax == 0) continue;
    dim3 grid((unsigned)std::min<int64_t>(cdiv(cdiv(nmax, 4), 256), 32), (unsigned)c);
    gather_segments_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(s, dst);
    SCVI_LAUNCH_CHECK("scvi_gather_segments");
  }
  return SCVI_OK;
}

int scvi_adam_step_graph(float* p, const float* g, float* m, float* v, int64_t n, int64_t* step_counter, float lr,
                         float beta1, float beta2, float eps, float weight_decay, float grad_scale, void* stream) {
  using namespace scvi;
  SCVI_REQUIRE(step_counter, "scvi_adam_step_graph: NULL step counter");
  return adam_launch(p, g, m, v, n, step_counter, 0, lr, beta1, beta2, eps, weight_decay, grad_scale, stream, "scvi_adam_step_graph");
}

int scvi_elbo_reduce(const float* reconst, const float* kl, int64_t B, float kl_weight, const float* kl_weight_dev,
                     float* out3, void* stream) {
  using namespace scvi;
  SCVI_REQUIRE(reconst && kl && out3 && B > 0, "scvi_elbo_reduce: bad argument");
  elbo_reduce_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(reconst, kl, B, kl_weight, kl_weight_dev, out3);
  SCVI_LAUNCH_CHECK("scvi_elbo_reduce");
  return SCVI_OK;
}

int scvi_elbo_seed(int64_t B, float kl_weight, const float* kl_weight_dev, float scale, float* g_reconst, float* g_kl,
                   void* stream) {
  using namespace scvi;
  SCVI_REQUIRE(g_reconst && g_kl && B > 0, "scvi_elbo_seed: bad argument");
  elbo_seed_kernel<<<(unsigned)cdiv(B, 256), 256, 0, (cudaStream_t)stream>>>(B, kl_weight, kl_weight_dev, scale, g_reconst, g_kl);
  SCVI_LAUNCH_CHECK("scvi_elbo_seed");
  return SCVI_OK;
}
}
