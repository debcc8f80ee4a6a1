// Micro-benchmark: how fast can one SM pull random 8000-byte-pitch row segments into shared memory?
//   mode 0: 1-D bulk copies (TMA engine), `lanes` issuing lanes of one warp, S bytes per copy
//   mode 1: loader warps, ld.global.v4 -> st.shared (nw warps, U loads in flight per thread)
//   mode 2: loader warps, cp.async 16 B (LDGSTS, no registers): one warp instruction = one row segment of
//           512 B (S = 512) -- completion through cp.async.mbarrier.arrive.noinc on the stage's mbarrier
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o ubench_gather ubench_gather.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include <vector>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, uint32_t c) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(c));
}
__device__ __forceinline__ void mbar_expect(uint64_t* b, uint32_t tx) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(tx) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t ph) {
  asm volatile(
      "{\n.reg .pred p;\nW: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D;\nbra W;\nD:\n}" ::"r"(smem_u32(b)),
      "r"(ph)
      : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

// rows: [nrows] random row ids; each CTA walks its own slice.  One "batch" = 128 rows x S bytes.
template <int MODE>
__global__ void __launch_bounds__(1024, 1) k(const float* __restrict__ x, int64_t ldx, const int* __restrict__ rows,
                                             int nbatch, int S, int pitch, int lanes, int nw, int ncols_chunks,
                                             float* sink) {
  extern __shared__ __align__(128) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem);
  uint8_t* buf = smem + 128;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int stage_bytes = 128 * pitch;
  if (threadIdx.x == 0) {
    mbar_init(&bars[0], MODE == 2 ? nw * 32 : 1);
    mbar_init(&bars[1], MODE == 2 ? nw * 32 : 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const int* myrows = rows + (size_t)blockIdx.x * 128;
  if (MODE == 0) {
    if (warp == 0) {
      for (int b = 0; b < nbatch; ++b) {
        const int s = b & 1;
        // wait for the previous use of this stage to have landed (no consumer: just the copy itself)
        if (b >= 2) mbar_wait(&bars[s], ((b >> 1) - 1) & 1);
        if (lane == 0) mbar_expect(&bars[s], 128u * S);
        __syncwarp();
        const int64_t col = (int64_t)(b % ncols_chunks) * (S / 4);
        if (lanes == 32) {
          for (int r = lane; r < 128; r += 32)
            bulk_g2s(buf + s * stage_bytes + r * pitch, x + (int64_t)myrows[r] * ldx + col, S, &bars[s]);
        } else if (lane == 0) {
          for (int r = 0; r < 128; ++r)
            bulk_g2s(buf + s * stage_bytes + r * pitch, x + (int64_t)myrows[r] * ldx + col, S, &bars[s]);
        }
      }
      if (nbatch >= 1) mbar_wait(&bars[(nbatch - 1) & 1], ((nbatch - 1) >> 1) & 1);
      if (nbatch >= 2) mbar_wait(&bars[(nbatch - 2) & 1], ((nbatch - 2) >> 1) & 1);
    }
  } else if (MODE == 2) {
    if (warp < nw) {
      const int segs = S / 512;  // 512-byte pieces per row segment
      for (int b = 0; b < nbatch; ++b) {
        const int s = b & 1;
        if (b >= 2) mbar_wait(&bars[s], ((b >> 1) - 1) & 1);
        const int64_t col = (int64_t)(b % ncols_chunks) * (S / 4);
        for (int r = warp; r < 128; r += nw) {
          const float* src = x + (int64_t)myrows[r] * ldx + col + lane * 4;
          const uint32_t dst = smem_u32(buf + s * stage_bytes + r * pitch + lane * 16);
          for (int g = 0; g < segs; ++g)
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + g * 512), "l"(src + g * 128) : "memory");
        }
        asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_u32(&bars[s])) : "memory");
      }
      if (nbatch >= 1) mbar_wait(&bars[(nbatch - 1) & 1], ((nbatch - 1) >> 1) & 1);
      if (nbatch >= 2) mbar_wait(&bars[(nbatch - 2) & 1], ((nbatch - 2) >> 1) & 1);
    }
  } else {
    // loader warps: batch = 128 rows x S bytes = 128*S/16 float4; thread t handles float4 #t, +nthreads ...
    if (warp < nw) {
      const int nth = nw * 32, t = threadIdx.x;
      const int per_row = S / 16;
      float acc = 0.f;
      for (int b = 0; b < nbatch; ++b) {
        const int s = b & 1;
        const int64_t col = (int64_t)(b % ncols_chunks) * (S / 4);
        const int total = 128 * per_row;
        for (int base = 0; base < total; base += nth * 8) {
          float4 v[8];
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            const int e = base + u * nth + t;
            if (e < total) {
              const int r = e / per_row, c = e % per_row;
              v[u] = __ldcs(reinterpret_cast<const float4*>(x + (int64_t)myrows[r] * ldx + col) + c);
            }
          }
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            const int e = base + u * nth + t;
            if (e < total) {
              const int r = e / per_row, c = e % per_row;
              *reinterpret_cast<float4*>(buf + s * stage_bytes + r * pitch + c * 16) = v[u];
            }
          }
        }
      }
      if (acc == 123.f) sink[0] = acc;
    }
  }
  __syncthreads();
  if (threadIdx.x == 0 && sink) sink[blockIdx.x + 1] = *reinterpret_cast<float*>(buf + 4 * (blockIdx.x % 32));
}

int main() {
  const int64_t nrows = 1 << 19, G = 2000;  // 4.2 GB
  float* x;
  cudaMalloc(&x, nrows * G * 4);
  cudaMemset(x, 0, nrows * G * 4);
  const int nsm = 148;
  std::vector<int> h(nsm * 128);
  srand(1);
  for (auto& v : h) v = (int)(((int64_t)rand() * 65536 + rand()) % nrows);
  int* rows;
  cudaMalloc(&rows, h.size() * 4);
  cudaMemcpy(rows, h.data(), h.size() * 4, cudaMemcpyHostToDevice);
  float* sink;
  cudaMalloc(&sink, 4096);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  const int smem_max = 200 * 1024;
  cudaFuncSetAttribute(k<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_max);
  cudaFuncSetAttribute(k<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_max);
  cudaFuncSetAttribute(k<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_max);
  auto run = [&](int mode, int S, int pad, int lanes, int nw) {
    const int pitch = S + pad;
    const int ncc = (int)(G * 4 / S);
    const int nbatch = 16 * 512 / S * 4;  // same bytes for every S: 128 rows * 32 KB
    const size_t smem = 128 + 2 * 128 * (size_t)pitch;
    if (smem > smem_max) return;
    float best = 1e9;
    for (int rep = 0; rep < 5; ++rep) {
      cudaEventRecord(e0);
      if (mode == 0) k<0><<<nsm, 1024, smem>>>(x, G, rows, nbatch, S, pitch, lanes, nw, ncc, sink);
      else if (mode == 1) k<1><<<nsm, 1024, smem>>>(x, G, rows, nbatch, S, pitch, lanes, nw, ncc, sink);
      else k<2><<<nsm, 1024, smem>>>(x, G, rows, nbatch, S, pitch, lanes, nw, ncc, sink);
      cudaEventRecord(e1);
      cudaEventSynchronize(e1);
      float ms;
      cudaEventElapsedTime(&ms, e0, e1);
      if (ms < best) best = ms;
    }
    cudaError_t err = cudaGetLastError();
    const double bytes = (double)nsm * nbatch * 128 * S;
    printf("mode=%d S=%4d pad=%2d lanes=%2d nw=%2d : %8.3f ms  %7.1f GB/s  (%.1f M copies/s/SM) %s\n", mode, S, pad, lanes, nw,
           best, bytes / best / 1e6, mode == 0 ? (double)nbatch * 128 / best / 1e3 : 0.0,
           err == cudaSuccess ? "" : cudaGetErrorString(err));
  };
  for (int S : {128, 256, 512, 1024}) {
    run(0, S, 16, 32, 0);
    run(0, S, 0, 32, 0);
    run(0, S, 16, 1, 0);
  }
  for (int S : {128, 512}) {
    for (int nw : {4, 8, 16, 32}) run(1, S, 16, 0, nw);
  }
  for (int nw : {1, 2, 4, 8}) run(2, 512, 16, 0, nw);
  if (getenv("UBENCH_1K")) for (int nw : {2, 4}) run(2, 1024, 16, 0, nw);
  return 0;
}
