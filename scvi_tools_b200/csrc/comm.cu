// Data-parallel gradient exchange fused with the optimizer: one kernel all-reduces a slice of the flat gradient
// arena over NVLink peer memory (one-shot: every rank reads every rank's slice, fixed rank order, so all
// replicas compute bit-identical sums) and applies the Adam update to it.
//
// Replaces torch DDP's bucketed ncclAllReduce + torch.optim.Adam.step for the SCVI training plan (reference
// src/scvi/train/_trainingplans.py:475-519 under `Trainer(strategy="ddp")`, tests/model/test_multigpu.py:29-36).
//
// Peer memory: each rank cudaMalloc's one "exchange segment" {gradient arena x 2 (step parity), flags, sequence
// counters}, exports it with cudaIpcGetMemHandle, and opens the other ranks' handles (single node, NVLink /
// NVSwitch).  Protocol of one launch on slot s (s = bucket id):
//   1. seq = ++sequence[s] (identical on all ranks: every rank runs the same launches in the same order);
//      block 0 writes flags[s][my rank] = seq into EVERY rank's segment (st.release.sys after a system fence:
//      the gradient kernels before this launch on the stream have completed, their writes are in this GPU's L2)
//   2. every block waits until its own segment shows flags[s][r] >= seq for all r (ld.acquire.sys)
//   3. sum_r g_r[parity][i] in rank order (ld.relaxed.sys: never a stale L1 line), Adam, write p / m / v locally
// Re-use of a gradient buffer is safe without a second barrier because the arena is double-buffered by step
// parity: rank A overwrites g[parity] two steps later, after it has passed the barrier of the step in between,
// which every peer only signals once its reads of the earlier step are complete (stream order).
// Waits are bounded: a peer that never arrives traps the kernel (error return) instead of hanging the GPU.
#include <string.h>

#include <algorithm>

#include "common.cuh"

namespace scvi {

constexpr int COMM_MAX_PEERS = 16;
constexpr int COMM_SLOTS = 4;
// layout of the control block at the end of a segment (uint64 units)
constexpr int COMM_FLAGS = 0;                             // [COMM_SLOTS][COMM_MAX_PEERS]
constexpr int COMM_SEQ = COMM_SLOTS * COMM_MAX_PEERS;     // [COMM_SLOTS][2] = {launches done, arrival ticket}
constexpr int COMM_CTRL_WORDS = COMM_SEQ + 2 * COMM_SLOTS;

struct PeerTable {
  const float* g[COMM_MAX_PEERS];
  unsigned long long* ctrl[COMM_MAX_PEERS];
};

__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 ld_relaxed_sys4(const float* p) {
  float4 r;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p) : "memory");
  return r;
}
__device__ __forceinline__ float ld_relaxed_sys(const float* p) {
  float r;
  asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(r) : "l"(p) : "memory");
  return r;
}

__device__ __forceinline__ void adam_update(float& p, float g, float& m, float& v, float grad_scale, float wd,
                                            float beta1, float beta2, float step_size, float bc2_sqrt, float eps) {
  g = g * grad_scale;  // same op order as adam.cu::adam_elem (torch's single-tensor Adam)
  g = fmaf(wd, p, g);
  m = m + (g - m) * (1.f - beta1);
  v = v * beta2 + (1.f - beta2) * g * g;
  const float denom = sqrtf(v) / bc2_sqrt + eps;
  p = p - step_size * (m / denom);
}

__global__ void __launch_bounds__(256)
    allreduce_adam_kernel(PeerTable pt, int world, int rank, int slot, float* __restrict__ p, float* __restrict__ m,
                          float* __restrict__ v, float* __restrict__ g_sum /*nullable: reduced gradient out*/,
                          int64_t begin, int64_t end, int64_t* step_counter, int advance, float lr, float beta1,
                          float beta2, float eps, float wd, float grad_scale) {
  __shared__ int s_fail;
  unsigned long long* my = pt.ctrl[rank];
  const unsigned long long seq = *reinterpret_cast<volatile unsigned long long*>(my + COMM_SEQ + 2 * slot) + 1ull;
  if (threadIdx.x == 0) s_fail = 0;
  // ---- 1. announce: this rank's gradients of the step are complete
  if (blockIdx.x == 0 && threadIdx.x < world) {
    __threadfence_system();
    st_release_sys(pt.ctrl[threadIdx.x] + COMM_FLAGS + slot * COMM_MAX_PEERS + rank, seq);
  }
  __syncthreads();
  // ---- 2. wait for every rank's announcement (flags live in this rank's own memory)
  if (threadIdx.x < world) {
    const unsigned long long* f = my + COMM_FLAGS + slot * COMM_MAX_PEERS + threadIdx.x;
    unsigned spins = 0;
    while (ld_acquire_sys(f) < seq) {
      __nanosleep(32);
      if (++spins > (1u << 26)) {  // several seconds: a peer died or ran a different launch sequence
        s_fail = 1;
        break;
      }
    }
  }
  __syncthreads();
  if (s_fail) {
    if (threadIdx.x == 0 && blockIdx.x == 0)
      printf("scvi_b200: gradient exchange timed out (rank %d, slot %d, launch %llu)\n", rank, slot, seq);
    __trap();
  }
  // ---- 3. reduce in rank order + Adam
  const int64_t step = *reinterpret_cast<volatile int64_t*>(step_counter) + 1;
  const double bc1 = 1.0 - pow((double)beta1, (double)step);
  const double bc2 = 1.0 - pow((double)beta2, (double)step);
  const float step_size = (float)((double)lr / bc1);
  const float bc2_sqrt = (float)sqrt(bc2);
  const int64_t n4 = (end - begin) >> 2;  // begin and end are multiples of 4 (16-byte aligned slices)
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    const int64_t e = begin + 4 * i;
    float4 part[COMM_MAX_PEERS];
#pragma unroll
    for (int r = 0; r < COMM_MAX_PEERS; ++r)
      if (r < world) part[r] = ld_relaxed_sys4(pt.g[r] + e);
    float4 gg = part[0];
#pragma unroll
    for (int r = 1; r < COMM_MAX_PEERS; ++r)
      if (r < world) {
        gg.x += part[r].x; gg.y += part[r].y; gg.z += part[r].z; gg.w += part[r].w;
      }
    float4 pp = *reinterpret_cast<float4*>(p + e);
    float4 mm = *reinterpret_cast<float4*>(m + e);
    float4 vv = *reinterpret_cast<float4*>(v + e);
    if (g_sum) *reinterpret_cast<float4*>(g_sum + e) = gg;
    adam_update(pp.x, gg.x, mm.x, vv.x, grad_scale, wd, beta1, beta2, step_size, bc2_sqrt, eps);
    adam_update(pp.y, gg.y, mm.y, vv.y, grad_scale, wd, beta1, beta2, step_size, bc2_sqrt, eps);
    adam_update(pp.z, gg.z, mm.z, vv.z, grad_scale, wd, beta1, beta2, step_size, bc2_sqrt, eps);
    adam_update(pp.w, gg.w, mm.w, vv.w, grad_scale, wd, beta1, beta2, step_size, bc2_sqrt, eps);
    *reinterpret_cast<float4*>(p + e) = pp;
    *reinterpret_cast<float4*>(m + e) = mm;
    *reinterpret_cast<float4*>(v + e) = vv;
  }
  // ---- bookkeeping by the last block: launch sequence of the slot, and (tail bucket) Adam's step counter
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    unsigned long long* ticket = my + COMM_SEQ + 2 * slot + 1;
    if (atomicAdd(ticket, 1ull) == (unsigned long long)gridDim.x - 1) {
      *ticket = 0ull;
      my[COMM_SEQ + 2 * slot] = seq;
      if (advance) step_counter[0] = step;
    }
  }
}

// ---------------------------------------------------------------------------------------
// Sharded variant (reduce-scatter + Adam + all-gather of the PARAMETERS in one launch): rank r owns the r-th
// slice of the arena, sums that slice over all ranks' gradient buffers (rank order), applies Adam to it with
// its own moments, and writes the updated parameters into EVERY rank's parameter arena (which therefore lives
// in the peer-mapped segment too).  Per rank and step: (N-1)/N of the arena read and as much written over
// NVLink, against (N-1) whole arenas read by the one-shot kernel above -- 2.7 MB each way instead of 21.7 MB at
// N = 8 -- and 1/N of the Adam work.  Every parameter is produced by exactly one rank with the arithmetic of the
// single-process step, so the replicas stay bit-identical by construction.  The moments of the other slices are
// not maintained on this rank.
// Protocol (slots s = arrival, s + 1 = completion):
//   1. arrival barrier as above: every rank has finished its backward, so nobody still READS parameters;
//   2. own slice: reduce, Adam, store p to all ranks;
//   3. the last block to finish fences (system scope), signals completion to every rank and waits for every
//      rank's completion signal: when the kernel ends, this rank's parameter arena holds every slice.
// ---------------------------------------------------------------------------------------
struct PeerParams {
  float* p[COMM_MAX_PEERS];
};

__global__ void __launch_bounds__(256)
    allreduce_adam_sharded_kernel(PeerTable pt, PeerParams pp, int world, int rank, int slot, float* __restrict__ m,
                                  float* __restrict__ v, int64_t n, int64_t* step_counter, float lr, float beta1,
                                  float beta2, float eps, float wd, float grad_scale) {
  __shared__ int s_fail;
  unsigned long long* my = pt.ctrl[rank];
  const unsigned long long seq = *reinterpret_cast<volatile unsigned long long*>(my + COMM_SEQ + 2 * slot) + 1ull;
  if (threadIdx.x == 0) s_fail = 0;
  if (blockIdx.x == 0 && threadIdx.x < world) {
    __threadfence_system();
    st_release_sys(pt.ctrl[threadIdx.x] + COMM_FLAGS + slot * COMM_MAX_PEERS + rank, seq);
  }
  __syncthreads();
  if (threadIdx.x < world) {
    const unsigned long long* f = my + COMM_FLAGS + slot * COMM_MAX_PEERS + threadIdx.x;
    unsigned spins = 0;
    while (ld_acquire_sys(f) < seq) {
      __nanosleep(32);
      if (++spins > (1u << 26)) {
        s_fail = 1;
        break;
      }
    }
  }
  __syncthreads();
  if (s_fail) {
    if (threadIdx.x == 0 && blockIdx.x == 0)
      printf("scvi_b200: gradient exchange timed out (rank %d, slot %d, launch %llu)\n", rank, slot, seq);
    __trap();
  }
  const int64_t step = *reinterpret_cast<volatile int64_t*>(step_counter) + 1;
  const double bc1 = 1.0 - pow((double)beta1, (double)step);
  const double bc2 = 1.0 - pow((double)beta2, (double)step);
  const float step_size = (float)((double)lr / bc1);
  const float bc2_sqrt = (float)sqrt(bc2);
  const int64_t n4 = n >> 2;  // the arena is padded to a multiple of 4 floats
  const int64_t chunk = (n4 + world - 1) / world;
  const int64_t lo = min((int64_t)rank * chunk, n4), hi = min(lo + chunk, n4);
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  float* p_own = pp.p[rank];
  for (int64_t i = lo + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += stride) {
    const int64_t e = 4 * i;
    float4 part[COMM_MAX_PEERS];
#pragma unroll
    for (int r = 0; r < COMM_MAX_PEERS; ++r)
      if (r < world) part[r] = ld_relaxed_sys4(pt.g[r] + e);
    float4 gg = part[0];
#pragma unroll
    for (int r = 1; r < COMM_MAX_PEERS; ++r)
      if (r < world) {
        gg.x += part[r].x; gg.y += part[r].y; gg.z += part[r].z; gg.w += part[r].w;
      }
    float4 q = *reinterpret_cast<float4*>(p_own + e);
    float4 mm = *reinterpret_cast<float4*>(m + e);
    float4 vv = *reinterpret_cast<float4*>(v + e);
    adam_update(q.x, gg.x, mm.x, vv.x, grad_scale, wd, beta1, beta2, step_size, bc2_sqrt, eps);
    adam_update(q.y, gg.y, mm.y, vv.y, grad_scale, wd, beta1, beta2, step_size, bc2_sqrt, eps);
    adam_update(q.z, gg.z, mm.z, vv.z, grad_scale, wd, beta1, beta2, step_size, bc2_sqrt, eps);
    adam_update(q.w, gg.w, mm.w, vv.w, grad_scale, wd, beta1, beta2, step_size, bc2_sqrt, eps);
    *reinterpret_cast<float4*>(m + e) = mm;
    *reinterpret_cast<float4*>(v + e) = vv;
#pragma unroll
    for (int r = 0; r < COMM_MAX_PEERS; ++r)
      if (r < world) *reinterpret_cast<float4*>(pp.p[r] + e) = q;
  }
  // ---- completion: last block of this rank signals, then waits for every rank's signal
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();  // this block's parameter stores are visible system-wide before the ticket
    unsigned long long* ticket = my + COMM_SEQ + 2 * slot + 1;
    if (atomicAdd(ticket, 1ull) == (unsigned long long)gridDim.x - 1) {
      __threadfence_system();
      for (int r = 0; r < world; ++r)
        st_release_sys(pt.ctrl[r] + COMM_FLAGS + (slot + 1) * COMM_MAX_PEERS + rank, seq);
      bool ok = true;
      for (int r = 0; r < world && ok; ++r) {
        const unsigned long long* f = my + COMM_FLAGS + (slot + 1) * COMM_MAX_PEERS + r;
        unsigned spins = 0;
        while (ld_acquire_sys(f) < seq) {
          __nanosleep(32);
          if (++spins > (1u << 26)) {
            ok = false;
            break;
          }
        }
      }
      *ticket = 0ull;
      my[COMM_SEQ + 2 * slot] = seq;
      step_counter[0] = step;
      if (!ok) {
        printf("scvi_b200: parameter exchange timed out (rank %d, launch %llu)\n", rank, seq);
        __trap();
      }
    }
  }
}

}  // namespace scvi

extern "C" {

int64_t scvi_comm_ctrl_bytes(void) { return (int64_t)scvi::COMM_CTRL_WORDS * 8; }

int scvi_comm_alloc(int64_t bytes, void** dev_ptr, void* ipc_handle64) {
  using namespace scvi;
  SCVI_REQUIRE(bytes > 0 && dev_ptr && ipc_handle64, "scvi_comm_alloc: bad argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  void* p = nullptr;
  cudaError_t e = cudaMalloc(&p, (size_t)bytes);
  if (e == cudaSuccess) e = cudaMemset(p, 0, (size_t)bytes);
  cudaIpcMemHandle_t h;
  if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, p);
  if (e != cudaSuccess) {
    if (p) cudaFree(p);
    cudaGetLastError();
    set_error("scvi_comm_alloc: %s", cudaGetErrorString(e));
    return SCVI_ECUDA;
  }
  memcpy(ipc_handle64, &h, 64);
  *dev_ptr = p;
  return SCVI_OK;
}

int scvi_comm_open(const void* ipc_handle64, void** dev_ptr) {
  using namespace scvi;
  SCVI_REQUIRE(ipc_handle64 && dev_ptr, "scvi_comm_open: bad argument");
  cudaIpcMemHandle_t h;
  memcpy(&h, ipc_handle64, 64);
  void* p = nullptr;
  cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
  if (e != cudaSuccess) {
    cudaGetLastError();
    set_error("scvi_comm_open: %s", cudaGetErrorString(e));
    return SCVI_ECUDA;
  }
  *dev_ptr = p;
  return SCVI_OK;
}

int scvi_comm_close(void* dev_ptr) {
  using namespace scvi;
  if (!dev_ptr) return SCVI_OK;
  cudaError_t e = cudaIpcCloseMemHandle(dev_ptr);
  if (e != cudaSuccess) {
    cudaGetLastError();
    set_error("scvi_comm_close: %s", cudaGetErrorString(e));
    return SCVI_ECUDA;
  }
  return SCVI_OK;
}

int scvi_comm_free(void* dev_ptr) {
  if (dev_ptr) cudaFree(dev_ptr);
  return SCVI_OK;
}

int scvi_allreduce_adam(const void* const* peer_g, void* const* peer_ctrl, int world, int rank, int slot, float* p,
                        float* m, float* v, float* g_sum, int64_t begin, int64_t end, int64_t* step_counter,
                        int advance_counter, float lr, float beta1, float beta2, float eps, float weight_decay,
                        int blocks, void* stream) {
  using namespace scvi;
  SCVI_REQUIRE(peer_g && peer_ctrl && world >= 1 && world <= COMM_MAX_PEERS && rank >= 0 && rank < world,
               "scvi_allreduce_adam: world must be 1..%d", COMM_MAX_PEERS);
  SCVI_REQUIRE(slot >= 0 && slot < COMM_SLOTS, "scvi_allreduce_adam: slot must be 0..%d", COMM_SLOTS - 1);
  SCVI_REQUIRE(p && m && v && step_counter && begin >= 0 && end >= begin, "scvi_allreduce_adam: bad argument");
  SCVI_REQUIRE(begin % 4 == 0 && end % 4 == 0, "scvi_allreduce_adam: slice bounds must be multiples of 4 floats");
  PeerTable pt{};
  for (int r = 0; r < world; ++r) {
    SCVI_REQUIRE(peer_g[r] && peer_ctrl[r] && ((uintptr_t)peer_g[r] % 16 == 0), "scvi_allreduce_adam: bad peer pointer %d", r);
    pt.g[r] = (const float*)peer_g[r];
    pt.ctrl[r] = (unsigned long long*)peer_ctrl[r];
  }
  const int64_t n4 = (end - begin) / 4;
  int64_t nb = blocks > 0 ? blocks : 148 * 2;
  nb = std::max<int64_t>(1, std::min<int64_t>(nb, cdiv(std::max<int64_t>(n4, 1), 256)));
  allreduce_adam_kernel<<<(unsigned)nb, 256, 0, (cudaStream_t)stream>>>(pt, world, rank, slot, p, m, v, g_sum, begin, end,
                                                                      step_counter, advance_counter, lr, beta1, beta2,
                                                                      eps, weight_decay, 1.f / (float)world);
  SCVI_LAUNCH_CHECK("scvi_allreduce_adam");
  return SCVI_OK;
}

int scvi_allreduce_adam_sharded(const void* const* peer_g, void* const* peer_p, void* const* peer_ctrl, int world,
                                int rank, int slot, float* m, float* v, int64_t n, int64_t* step_counter, float lr,
                                float beta1, float beta2, float eps, float weight_decay, int blocks, void* stream) {
  using namespace scvi;
  SCVI_REQUIRE(peer_g && peer_p && peer_ctrl && world >= 1 && world <= COMM_MAX_PEERS && rank >= 0 && rank < world,
               "scvi_allreduce_adam_sharded: world must be 1..%d", COMM_MAX_PEERS);
  SCVI_REQUIRE(slot >= 0 && slot + 1 < COMM_SLOTS, "scvi_allreduce_adam_sharded: slot must be 0..%d", COMM_SLOTS - 2);
  SCVI_REQUIRE(m && v && step_counter && n >= 0 && n % 4 == 0,
               "scvi_allreduce_adam_sharded: bad argument (n must be a multiple of 4 floats)");
  PeerTable pt{};
  PeerParams pp{};
  for (int r = 0; r < world; ++r) {
    SCVI_REQUIRE(peer_g[r] && peer_p[r] && peer_ctrl[r] && ((uintptr_t)peer_g[r] % 16 == 0) && ((uintptr_t)peer_p[r] % 16 == 0),
                 "scvi_allreduce_adam_sharded: bad peer pointer %d", r);
    pt.g[r] = (const float*)peer_g[r];
    pt.ctrl[r] = (unsigned long long*)peer_ctrl[r];
    pp.p[r] = (float*)peer_p[r];
  }
  const int64_t chunk = cdiv(std::max<int64_t>(n / 4, 1), world);
  int64_t nb = blocks > 0 ? blocks : 148 * 2;
  nb = std::max<int64_t>(1, std::min<int64_t>(nb, cdiv(chunk, 256)));
  allreduce_adam_sharded_kernel<<<(unsigned)nb, 256, 0, (cudaStream_t)stream>>>(pt, pp, world, rank, slot, m, v, n, step_counter, lr,
                                                                               beta1, beta2, eps, weight_decay,
                                                                               1.f / (float)world);
  SCVI_LAUNCH_CHECK("scvi_allreduce_adam_sharded");
  return SCVI_OK;
}
}
