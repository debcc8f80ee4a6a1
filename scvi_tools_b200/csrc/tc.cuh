// sm_100a primitives used by the tensor-core kernels: mbarrier, 1-D bulk async copy (TMA engine,
// no tensor map), tcgen05 alloc / mma / commit / ld, shared-memory matrix descriptors.
// Hand-written inline PTX (no CUTLASS dependency).
#pragma once
#include <cuda_fp16.h>

#include "common.cuh"

namespace scvi {
namespace tc {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ---- mbarrier ---------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// The suspend-time hint (ns) lets the hardware park the thread until the phase completes instead of returning
// after its short default limit: a waiting role (producer, MMA issuer) then costs no issue slots on the
// scheduler it shares with the epilogue warps (ncu, r01i: 16 % of K3's executed instructions were this loop).
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred P1;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2, %3;\n\t"
      "selp.b32 %0, 1, 0, P1;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity), "r"(20000u)
      : "memory");
  return ok != 0;
}
// Bounded wait: a protocol bug traps (the launch fails with an error) instead of hanging the GPU.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    if (++spins > 4000000u) {
      printf("scvi_b200: mbarrier wait timed out (block %d thread %d)\n", blockIdx.x, threadIdx.x);
      __trap();
    }
  }
}

// ---- async proxy / bulk copy ------------------------------------------------------------------
// generic-proxy smem writes (st.shared) -> visible to the async proxy (tcgen05.mma / bulk copy)
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// global -> shared, `bytes` multiple of 16, both addresses 16-byte aligned; completes on `bar`
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(smem_dst)), "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

// 16-byte asynchronous copy global -> shared issued by a thread (LDGSTS, L2-only caching, no register
// staging); generic proxy, so ordinary shared loads may read the data once the tracking mbarrier flips
__device__ __forceinline__ void cp_async16(uint32_t smem_dst, const void* gmem_src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_dst), "l"(gmem_src) : "memory");
}
// `bar` receives one arrival (pre-counted in its init count: .noinc) once every cp.async this thread has
// issued so far has landed
__device__ __forceinline__ void cp_async_mbar_arrive_noinc(uint64_t* bar) {
  asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// ---- tcgen05 ----------------------------------------------------------------------------------
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_result, uint32_t ncols) {  // one full warp
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_result)), "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {  // the same warp
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void fence_before_sync() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_after_sync() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// D[tmem] (+)= A[smem desc] * B[smem desc]^T, kind::f16 (fp16/bf16 inputs, fp32 accumulate); one thread issues
__device__ __forceinline__ void mma_f16_ss(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                           uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// ---- MMA issue from a warp-uniform loop ---------------------------------------------------------------
// The issuing WARP runs its loop converged (all 32 lanes wait on the mbarriers), one lane elected once issues
// the MMAs inside `if (leader) { ... }`, and descriptors are passed as two 32-bit words whose low word only
// needs an integer add to advance the start address.  ptxas then keeps every operand on the uniform datapath:
// ~3 instructions per MMA (two UMOV + UTCHMMA).  Issued the old way -- inside an `if (lane == 0)` region with
// 64-bit descriptors rebuilt per MMA -- every MMA sat in its own ELECT / BRA.U.ANY waterfall loop behind ~8
// dependent uniform-pipe ops: ~70 cycles per MMA, which made the single issuing thread the critical path of the
// fused decoder kernels (66 MMAs per 32-gene tile ~ 4.3k cycles, measured 2.2 us per tile; r02c capture).
__device__ __forceinline__ uint32_t elect_one() {  // call with the warp converged; exactly one lane gets 1
  uint32_t pred = 0;
  asm volatile("{\n\t.reg .pred P1;\n\telect.sync _|P1, 0xffffffff;\n\tselp.b32 %0, 1, 0, P1;\n\t}" : "=r"(pred));
  return pred;
}
// descriptor words: lo = start address (>> 4, 14 bits) | LBO (>> 4) << 16;  hi = SBO (>> 4) | version 1 (bit 46)
__device__ __forceinline__ uint32_t desc_lo(uint32_t smem_addr, uint32_t lbo_bytes) {
  return ((smem_addr >> 4) & 0x3FFFu) | (((lbo_bytes >> 4) & 0x3FFFu) << 16);
}
__host__ __device__ constexpr uint32_t desc_hi(uint32_t sbo_bytes) { return ((sbo_bytes >> 4) & 0x3FFFu) | (1u << 14); }
__device__ __forceinline__ void mma_f16_ss2(uint32_t d_tmem, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo, uint32_t b_hi,
                                            uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\t"
      "setp.ne.b32 p, %6, 0;\n\t"
      "mov.b64 da, {%1, %2};\n\t"
      "mov.b64 db, {%3, %4};\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %5, p;\n\t}"
      ::"r"(d_tmem), "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accumulate)
      : "memory");
}
// One operand of the error-compensated product: its fp16 `hi` and `lo` blocks share LBO / SBO and the K-step.
struct Oper {
  uint32_t hi_lo, lo_lo;  // low descriptor words of the hi block and of the lo block
  uint32_t dhi;           // high descriptor word (SBO, version)
  uint32_t step;          // advance of the start-address field per K-step of 16 (bytes >> 4)
};
__device__ __forceinline__ Oper make_oper(uint32_t hi_addr, uint32_t lo_addr, uint32_t lbo, uint32_t sbo, uint32_t step_bytes) {
  Oper o;
  o.hi_lo = desc_lo(hi_addr, lbo);
  o.lo_lo = desc_lo(lo_addr, lbo);
  o.dhi = desc_hi(sbo);
  o.step = step_bytes >> 4;
  return o;
}
// D (+)= A * B over KS K-steps of 16 with the 3-MMA split  hi*hi + hi*lo + lo*hi.  Call inside `if (leader)`.
template <int KS>
__device__ __forceinline__ void mma3(uint32_t d_tmem, const Oper& a, const Oper& b, uint32_t idesc, bool accumulate) {
#pragma unroll
  for (int ks = 0; ks < KS; ++ks) {
    const uint32_t ah = a.hi_lo + ks * a.step, al = a.lo_lo + ks * a.step;
    const uint32_t bh = b.hi_lo + ks * b.step, bl = b.lo_lo + ks * b.step;
    mma_f16_ss2(d_tmem, ah, a.dhi, bh, b.dhi, idesc, (accumulate || ks > 0) ? 1u : 0u);
    mma_f16_ss2(d_tmem, ah, a.dhi, bl, b.dhi, idesc, 1u);
    mma_f16_ss2(d_tmem, al, a.dhi, bh, b.dhi, idesc, 1u);
  }
}

// mbarrier arrives when all tcgen05 ops issued so far by this thread have completed
__device__ __forceinline__ void mma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}

// TMEM -> registers: lane i of the warp gets 16 consecutive fp32 columns of TMEM lane (32*(warp%4) + i)
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16]) {
  uint32_t r[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr));
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, float (&v)[8]) {
  uint32_t r[8];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr));
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// ---- descriptors ------------------------------------------------------------------------------
// K-major operand, no swizzle ("interleave"): element (row r, 16-byte k-chunk c) at
//   base + (r % 8) * 16 + (r / 8) * SBO + c * LBO      (8 rows x 16 B core matrices, 128 B contiguous)
// Our tiles use  addr(r, c) = c * (ROWS * 16) + r * 16  =>  SBO = 128 B, LBO = ROWS * 16 B.
__device__ __forceinline__ uint64_t smem_desc_kmajor(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;  // descriptor version 1 (Blackwell)
  return d;                // base_offset 0, lbo_mode 0, layout_type 0 = SWIZZLE_NONE
}
// MN-major operand, no swizzle: element (mn index i, reduction index k) at
//   base + (k % 8) * 16 + (k / 8) * LBO + (i / 8) * SBO + (i % 8) * 2     (fp16)
// i.e. 8 reduction rows x 16 B (8 contiguous MN elements) core matrices.  The same descriptor encoding as
// the K-major one; the major-ness is selected by the instruction descriptor.
__device__ __forceinline__ uint64_t smem_desc(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  return smem_desc_kmajor(smem_addr, lbo_bytes, sbo_bytes);
}
// instruction descriptor, kind::f16: F16 x F16 -> F32, M x N tile; a_mn / b_mn = 1 selects an MN-major operand
__host__ __device__ constexpr uint32_t idesc_f16_f32(int M, int N, int a_mn = 0, int b_mn = 0) {
  return (1u << 4)                      // c_format = F32
         | (0u << 7) | (0u << 10)       // a_format = b_format = F16
         | ((uint32_t)a_mn << 15) | ((uint32_t)b_mn << 16)
         | ((uint32_t)(N >> 3) << 17)   // n_dim
         | ((uint32_t)(M >> 4) << 24);  // m_dim
}
// exact power of two 2^e as a float (|e| <= 126)
__device__ __forceinline__ float pow2i(int e) { return __int_as_float((127 + e) << 23); }
// 8 floats (already scaled into fp16 range) -> one 16-byte chunk of fp16 `hi` and one of `lo` residuals
__device__ __forceinline__ void split8(const float (&v)[8], uint4& hi, uint4& lo) {
  __half2 h[4], l[4];
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const __half2 hh = __floats2half2_rn(v[2 * q], v[2 * q + 1]);
    const float2 back = __half22float2(hh);
    h[q] = hh;
    l[q] = __floats2half2_rn(v[2 * q] - back.x, v[2 * q + 1] - back.y);
  }
  hi = *reinterpret_cast<uint4*>(h);
  lo = *reinterpret_cast<uint4*>(l);
}

// ---- fp16 hi/lo split: v ~= hi + lo with |err| <= 2^-22 |v| for |v| in fp16 normal range ---------
__device__ __forceinline__ void split_f16(float v, __half& hi, __half& lo) {
  hi = __float2half_rn(v);
  lo = __float2half_rn(v - __half2float(hi));
}

}  // namespace tc
}  // namespace scvi
