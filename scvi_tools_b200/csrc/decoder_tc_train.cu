// tcgen05 / TMEM training kernels of the fused decoder (SCVI_PREC_TC):
//
//   K2  tc_nll_dh_kernel   (CTA = 128 rows x a range of 32-gene tiles, "rows-owned")
//       logits GEMM -> per-element NB/ZINB likelihood AND its per-element derivatives in one pass ->
//       reconst partials, S = sum_g g1, and dh through three more tensor-core GEMMs whose A operands
//       (g1, softmax p, -dpi) are staged by the epilogue warps as fp16 hi/lo blocks:
//           dh_a = g1 * Ws + (-dpi) * Wp          dh_b = p * Ws          dh = u * (dh_a - S * dh_b)
//       (the softmax backward term needs the full-row S, hence the second accumulator).
//   K3  tc_dw_kernel       (CTA = 128 genes x a range of 32-row tiles, "genes-owned", transposed tile)
//       logits^T GEMM -> per-element derivatives -> ds^T, dp^T staged -> dWs, dWp accumulate in TMEM
//       over all the CTA's rows; column sums dbs, dbp, dpx_r in registers.  Partial slabs per row split,
//       reduced in fixed order (deterministic, no atomics).
//
// Replaces autograd's backward of DecoderSCVI.forward's heads + log_prob (reference
// src/scvi/nn/_base_components.py:436-441, src/scvi/distributions/_negative_binomial.py:36-135).
// Warp roles: warp 0 = bulk-copy producer, warp 1 = MMA issuer (one lane), warp 2 = TMEM allocator,
// warps 4..19 = epilogue (thread = one TMEM lane x 8 columns).
#include "decoder_math.cuh"
#include "tc_layout.cuh"

namespace scvi {

constexpr int TT_EPI_WARPS = 16;
constexpr int TT_THREADS = 32 * (4 + TT_EPI_WARPS);
constexpr uint32_t TT_ABLK = TC_M * TC_NT * 2;  // staged A block: 128 x 32 fp16 = 8 KB

// fp32 [rows, H] -> scaled fp16 hi / lo blocks (128-row block layout) in shared memory
__device__ __forceinline__ void tt_stage_h(const float* __restrict__ h, int64_t ldh, int valid, uint8_t* s_hi,
                                           uint8_t* s_lo, int tid, int nthreads) {
  for (int e = tid; e < TC_M * TC_KC; e += nthreads) {
    const int r = e % TC_M, kc = e / TC_M;
    float v[8];
    if (r < valid) {
      const float4 a = *reinterpret_cast<const float4*>(h + (int64_t)r * ldh + 8 * kc);
      const float4 b = *reinterpret_cast<const float4*>(h + (int64_t)r * ldh + 8 * kc + 4);
      v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
#pragma unroll
      for (int q = 0; q < 8; ++q) v[q] = fminf(fmaxf(v[q] * TC_SCALE_H, -60000.f), 60000.f);
    } else {
#pragma unroll
      for (int q = 0; q < 8; ++q) v[q] = 0.f;
    }
    uint4 hi, lo;
    tc::split8(v, hi, lo);
    *reinterpret_cast<uint4*>(s_hi + kc * TC_M * 16 + r * 16) = hi;
    *reinterpret_cast<uint4*>(s_lo + kc * TC_M * 16 + r * 16) = lo;
  }
}

// =======================================================================================
// Count operand of the training kernels.  XK selects how the resident matrix is read:
//   XK_F32      fp32 counts, any value (generic per-element path: table or lgamma series, per element)
//   XK_F32_INT  fp32 counts the caller guarantees to be non-negative integers (SCVI_X_INTEGER)
//   XK_U8       uint8 counts (SCVI_X_U8): 1 byte per count from HBM, no conversion instructions
// For the two integer kinds a thread checks its 8 counts of a tile at once ("all below the table length");
// if so the straight-line lean element runs, otherwise the generic one.
// =======================================================================================
enum { XK_F32 = 0, XK_F32_INT = 1, XK_U8 = 2 };

template <int XK>
struct XRaw {
  uint32_t w[XK == XK_U8 ? 2 : 8];
};

// this thread's 8 counts of one row starting at gene gbb (zero beyond G or for an invalid row)
template <int XK>
__device__ __forceinline__ void tt_load_x8(const void* __restrict__ xrow_v, int64_t gbb, int64_t G, bool rvalid,
                                           int xvec, XRaw<XK>& d) {
  if (XK == XK_U8) {
    const uint8_t* xrow = reinterpret_cast<const uint8_t*>(xrow_v);
    if (rvalid && xvec && gbb + 8 <= G) {
      asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(d.w[0]), "=r"(d.w[1]) : "l"(xrow + gbb));
    } else {
      d.w[0] = d.w[1] = 0u;
#pragma unroll
      for (int i = 0; i < 8; ++i)
        if (rvalid && gbb + i < G) d.w[i >> 2] |= (uint32_t)__ldg(xrow + gbb + i) << (8 * (i & 3));
    }
  } else {
    const float* xrow = reinterpret_cast<const float*>(xrow_v);
    if (rvalid && xvec) {
      const float4 t4 = (gbb < G) ? ld_stream4(xrow + gbb) : make_float4(0.f, 0.f, 0.f, 0.f);
      const float4 u4 = (gbb + 4 < G) ? ld_stream4(xrow + gbb + 4) : make_float4(0.f, 0.f, 0.f, 0.f);
      d.w[0] = __float_as_uint(t4.x); d.w[1] = __float_as_uint(t4.y); d.w[2] = __float_as_uint(t4.z); d.w[3] = __float_as_uint(t4.w);
      d.w[4] = __float_as_uint(u4.x); d.w[5] = __float_as_uint(u4.y); d.w[6] = __float_as_uint(u4.z); d.w[7] = __float_as_uint(u4.w);
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) d.w[i] = (rvalid && gbb + i < G) ? __float_as_uint(__ldg(xrow + gbb + i)) : 0u;
    }
  }
}

// raw -> counts as floats + table indices; returns "all 8 are integers below `limit`" (limit = 64: a power of two)
template <int XK>
__device__ __forceinline__ bool tt_prep_x8(const XRaw<XK>& d, float (&xq)[8], int (&xi)[8]) {
  constexpr float MAGIC = 8388608.f;  // 2^23: integer n < 2^23 -> float bits 0x4B000000 | n
  if (XK == XK_U8) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const uint32_t bits = __byte_perm(d.w[i >> 2], 0x4B000000u, 0x7650u | (uint32_t)(i & 3));
      xq[i] = __uint_as_float(bits) - MAGIC;
      xi[i] = (int)(bits & 0xffu);
    }
    return ((d.w[0] | d.w[1]) & 0xC0C0C0C0u) == 0u;
  } else if (XK == XK_F32_INT) {
    uint32_t orv = 0u;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      xq[i] = __uint_as_float(d.w[i]);
      const uint32_t bits = __float_as_uint(xq[i] + MAGIC);
      orv |= bits ^ 0x4B000000u;
      xi[i] = (int)(bits & 63u);
    }
    return (orv & ~63u) == 0u;
  } else {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      xq[i] = __uint_as_float(d.w[i]);
      xi[i] = 0;
    }
    return false;
  }
}
static_assert(fm::LGT == 64 && fm::DGT <= 64, "tt_prep_x8 tests counts against 64");

#ifdef SCVI_K2_TRACE
// Debug build only (make NVFLAGS+=-DSCVI_K2_TRACE): clock64 stamps of CTA (0,0) -- epilogue warp 4, epilogue warp 19
// and the MMA-issuing warp -- per tile, fetched with scvi_debug_k2_trace (tools/k2_trace.py).
__device__ long long g_k2_trace[3 * 64 * 8];
#define K2_TR(who, j, slot)                                                                      \
  do {                                                                                           \
    if (blockIdx.x == 0 && blockIdx.y == 0 && lane == 0 && (j) < 64) g_k2_trace[((who) * 64 + (j)) * 8 + (slot)] = clock64(); \
  } while (0)
#else
#define K2_TR(who, j, slot) do {} while (0)
#endif

// =======================================================================================
// K2: fused NLL + dh
// =======================================================================================
template <bool ZINB, bool AOPS = false>
struct K2Smem {
  static constexpr int NH = ZINB ? 2 : 1;
  static constexpr uint32_t WSTAGE = tc_w32_stage_bytes(NH);
  static constexpr int NMMA = ZINB ? 3 : 2;          // staged operands of the dh GEMMs: g1, p, (dp)
  static constexpr int NSTG = NMMA;  // (AOPS: the same blocks are also copied to global memory for the dW GEMM)
  static constexpr uint32_t H_HI = 0;
  static constexpr uint32_t H_LO = TC_OPER_BYTES;
  static constexpr uint32_t W0 = 2 * TC_OPER_BYTES;
  static constexpr uint32_t STG = W0 + 2 * WSTAGE;
  static constexpr uint32_t BARS = STG + NSTG * 2 * TT_ABLK;  // 11 mbarriers
  static constexpr uint32_t TMEM_PTR = BARS + 16 * 8;
  static constexpr uint32_t LGX1 = TMEM_PTR + 16;
  static constexpr uint32_t RED = LGX1 + fm::LGX1_TABLE * 4;  // [4 col groups][128 rows][2]
  static constexpr uint32_t TOTAL = RED + 4 * TC_M * 2 * 4;
};

// the 8 elements of one thread for one tile.  Padded genes (beyond G) carry constants that make every output an
// exact zero (scale bias -1e30 => mu = 0, dropout bias 0, theta 1, count 0), so nothing is masked here; rows
// beyond B produce values that are never stored.
template <bool ZINB, bool LEAN>
__device__ __forceinline__ void tt_k2_elems(const float (&xq)[8], const int (&xi)[8], const float (&sv)[8],
                                            const float (&pv)[8], const float* __restrict__ gc,
                                            const float* __restrict__ tab, const float* __restrict__ s_lgx1, float krow,
                                            float eps, float s1, float kP, float kD, float& nll, float& S,
                                            float (&a1)[8], float (&a2)[8], float (&a3)[8]) {
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const float4 k0 = *reinterpret_cast<const float4*>(gc + i * TC_GC);
    const float4 k1 = *reinterpret_cast<const float4*>(gc + i * TC_GC + 4);
    fm::GeneK gk;
    gk.theta = k0.x; gk.log_theta = k0.y; gk.lg_theta = k0.z; gk.rth = k1.y; gk.dg_theta = k1.z;
    const float sl = fmaf(sv[i], TC_RESCALE, k0.w);
    const float pl = ZINB ? fmaf(pv[i], TC_RESCALE, k1.x) : 0.f;
    float ll, g1, dpi, mu;
    if (LEAN)
      fm::elem_fwd_lean<ZINB>(xq[i], tab[i * fm::LGT + xi[i]], sl, pl, krow, gk, eps, ll, g1, dpi, mu);
    else
      fm::elem_fwd_tab<ZINB>(xq[i], sl, pl, krow, gk, eps, tab + i * fm::LGT, s_lgx1, ll, g1, dpi, mu);
    nll -= ll;
    S += g1;
    a1[i] = g1 * s1;
    a2[i] = mu * kP;
    a3[i] = dpi * kD;
  }
}

// Staged operands of K2 kept for the weight-gradient GEMM (SCVI_PREC_AOPS): part p of (row tile rt, 16-row piece o,
// 32-gene tile t, 8-gene chunk cg) = 16 rows x 16 B at
//   aops + p * tt_aops_part_bytes + ((((rt * 8 + o) * nT4 + t) * 4 + cg) * 256,    nT4 = tiles padded to whole 128-gene blocks
// i.e. the 4 KB block [16 gene chunks][16 rows][16 B] of (row piece, 128-gene block) is contiguous: an MN-major A
// operand of the GEMM, one bulk copy.  Parts: 0/1 = g1 s1 hi/lo, 2/3 = p 2^P hi/lo, 4/5 = -dpi 2^D hi/lo (ZINB).
__host__ __device__ inline int64_t tt_nT4(int64_t G) { return ((G + TC_M - 1) / TC_M) * 4; }
__host__ __device__ inline size_t tt_aops_part_bytes(int64_t B, int64_t G) {
  return (size_t)((B + TC_M - 1) / TC_M) * 8 * tt_nT4(G) * 4 * 256;
}
__host__ __device__ inline int tt_aops_parts(bool zinb) { return zinb ? 6 : 4; }
size_t tc_aops_bytes(int64_t B, int64_t G, bool zinb) { return tt_aops_part_bytes(B, G) * tt_aops_parts(zinb); }

template <bool ZINB, int XK, bool AOPS>
__global__ void __launch_bounds__(TT_THREADS, 1)
    tc_nll_dh_kernel(const float* __restrict__ h, int64_t ldh, int64_t B, int64_t G,
                     const uint8_t* __restrict__ packed32, const void* __restrict__ x, int64_t ldx,
                     const int64_t* __restrict__ row_idx, const float* __restrict__ library,
                     const float* __restrict__ lse_part /*[nsplit_lse][B][2]*/, int nsplit_lse, int tiles_per_split,
                     float eps, int xvec, float* __restrict__ out_part /*[nsplit][B][2] = (nll, S)*/,
                     float* __restrict__ dh_part /*[nsplit][2][B][H]*/, uint8_t* __restrict__ aops /*nullable*/) {
  using L = K2Smem<ZINB, AOPS>;
  constexpr bool KEEP = AOPS && XK != XK_F32;  // integer counts: the staged operands are kept for the dW GEMM
  constexpr int NH = L::NH;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((128u - (tc::smem_u32(smem_raw) & 127u)) & 127u);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::BARS);
  uint64_t *wfull = bars, *wempty = bars + 2, *tfull = bars + 4, *tempty = bars + 6;
  uint64_t *sfull = bars + 8, *sempty = bars + 9, *accfull = bars + 10;
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(smem + L::TMEM_PTR);
  float* s_lgx1 = reinterpret_cast<float*>(smem + L::LGX1);
  float* s_red = reinterpret_cast<float*>(smem + L::RED);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int64_t r0 = (int64_t)blockIdx.x * TC_M;
  const int valid = (int)min((int64_t)TC_M, B - r0);
  const int nT = (int)((G + TC_NT - 1) / TC_NT);
  const int t0 = blockIdx.y * tiles_per_split;
  const int t1 = min(nT, t0 + tiles_per_split);
  const int ntiles = max(0, t1 - t0);

  if (warp == 1 && lane == 0) {
    for (int i = 0; i < 2; ++i) {
      tc::mbar_init(&wfull[i], 1);
      tc::mbar_init(&wempty[i], 1);
      tc::mbar_init(&tfull[i], 1);
      tc::mbar_init(&tempty[i], TT_EPI_WARPS);
    }
    tc::mbar_init(sfull, TT_EPI_WARPS);
    tc::mbar_init(sempty, KEEP ? 3 : 1);  // + the two copier warps when the operands are kept
    tc::mbar_init(accfull, 1);
    tc::fence_mbar_init();
  }
  if (warp == 2) tc::tmem_alloc(s_tmem, 512);
  for (int i = tid; i < fm::LGX1_TABLE; i += TT_THREADS) s_lgx1[i] = lgammaf((float)i + 1.f);
  tt_stage_h(h + r0 * ldh, ldh, valid, smem + L::H_HI, smem + L::H_LO, tid, TT_THREADS);
  tc::fence_proxy_async_smem();
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem_base = *s_tmem;
  // TMEM columns: logits [buf][head] 32 each at 0..127; acc1 (g1*Ws) 128; acc2 (p*Ws) 256; acc3 (dp*Wp) 384
  const uint32_t ACC1 = tmem_base + 128, ACC2 = tmem_base + 256, ACC3 = tmem_base + 384;

  if (warp == 0) {
    // ===== producer: one 32-gene tile {Ws hi/lo, (Wp hi/lo), gene constants} per stage, one bulk copy
    if (lane == 0) {
      for (int j = 0; j < ntiles; ++j) {
        const int s = j & 1;
        tc::mbar_wait(&wempty[s], ((j >> 1) & 1) ^ 1);
        tc::mbar_arrive_expect_tx(&wfull[s], L::WSTAGE);
        tc::bulk_g2s(smem + L::W0 + s * L::WSTAGE, packed32 + (size_t)(t0 + j) * L::WSTAGE, L::WSTAGE, &wfull[s]);
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: the warp runs the loop converged, one elected lane issues (tc.cuh, "MMA issue from a
    // warp-uniform loop")
    if (ntiles > 0) {
      constexpr uint32_t idescL = tc::idesc_f16_f32(TC_M, TC_NT, 0, 0);  // logits: 128 x 32, both K-major
      constexpr uint32_t idescO = tc::idesc_f16_f32(TC_M, TC_H, 0, 1);   // dh: 128 x 128, B MN-major
      const uint32_t leader = tc::elect_one();
      const uint32_t stg = tc::smem_u32(smem + L::STG);
      const uint32_t w0 = tc::smem_u32(smem + L::W0);
      // A = h tile, K-major (128 rows x 128 hidden): LBO = 128*16, SBO = 128, K-step = 2 chunks
      const tc::Oper hA = tc::make_oper(tc::smem_u32(smem + L::H_HI), tc::smem_u32(smem + L::H_LO), TC_M * 16, 128, 2 * TC_M * 16);
      // staged A blocks (K-major, 128 rows x 32 genes): LBO = 128*16, SBO = 128, K-step = 2 chunks
      const tc::Oper sA1 = tc::make_oper(stg + 0 * TT_ABLK, stg + 1 * TT_ABLK, TC_M * 16, 128, 2 * TC_M * 16);
      const tc::Oper sA2 = tc::make_oper(stg + 2 * TT_ABLK, stg + 3 * TT_ABLK, TC_M * 16, 128, 2 * TC_M * 16);
      const tc::Oper sA3 = tc::make_oper(stg + 4 * TT_ABLK, stg + 5 * TT_ABLK, TC_M * 16, 128, 2 * TC_M * 16);
      // W blocks of weight stage 0 (stage 1 = + WSTAGE): as K-major B of the logits GEMM (N = 32 genes, K = hidden:
      // LBO = 32*16, SBO = 128, K-step = 2 chunks) and as MN-major B of the dh GEMMs (N = hidden, K = genes:
      // LBO = 128 (next 8 genes), SBO = 32*16, K-step = 256 B)
      tc::Oper wL[NH], wO[NH];
#pragma unroll
      for (int hd = 0; hd < NH; ++hd) {
        wL[hd] = tc::make_oper(w0 + hd * 2 * TC_BLK32, w0 + hd * 2 * TC_BLK32 + TC_BLK32, TC_NT * 16, 128, 2 * TC_NT * 16);
        wO[hd] = tc::make_oper(w0 + hd * 2 * TC_BLK32, w0 + hd * 2 * TC_BLK32 + TC_BLK32, 128, TC_NT * 16, 256);
      }
      auto stage_of = [](const tc::Oper& o, int s) {  // the same blocks in weight stage s
        tc::Oper r = o;
        r.hi_lo += s * (L::WSTAGE >> 4);
        r.lo_lo += s * (L::WSTAGE >> 4);
        return r;
      };
      auto issue_logits = [&](int j) {
        const int s = j & 1;
        tc::mbar_wait(&wfull[s], (j >> 1) & 1);
        tc::mbar_wait(&tempty[s], ((j >> 1) & 1) ^ 1);
        tc::fence_after_sync();
        if (leader) {
#pragma unroll
          for (int hd = 0; hd < NH; ++hd)
            tc::mma3<TC_H / 16>(tmem_base + (s * 2 + hd) * TC_NT, hA, stage_of(wL[hd], s), idescL, false);
          tc::mma_commit(&tfull[s]);
        }
        __syncwarp();
      };
      issue_logits(0);
      for (int j = 0; j < ntiles; ++j) {
        if (j + 1 < ntiles) issue_logits(j + 1);
        const int s = j & 1;
        K2_TR(2, j, 0);
        tc::mbar_wait(sfull, j & 1);
        K2_TR(2, j, 1);
        tc::fence_after_sync();
        if (leader) {
          tc::mma3<TC_NT / 16>(ACC1, sA1, stage_of(wO[0], s), idescO, j > 0);
          tc::mma3<TC_NT / 16>(ACC2, sA2, stage_of(wO[0], s), idescO, j > 0);
          if (ZINB) tc::mma3<TC_NT / 16>(ACC3, sA3, stage_of(wO[NH - 1], s), idescO, j > 0);
          tc::mma_commit(sempty);      // staging buffers reusable
          tc::mma_commit(&wempty[s]);  // weight stage reusable
        }
        __syncwarp();
        K2_TR(2, j, 2);
      }
      if (leader) tc::mma_commit(accfull);
      __syncwarp();
    }
  } else if ((warp == 2 || warp == 3) && KEEP) {
    // ===== copier (SCVI_PREC_AOPS): the staged operand blocks of every tile go to global memory in the layout the dW
    // GEMM loads (tt_aops_part_bytes).  A whole warp that does nothing else: the epilogue warps pay nothing for it
    // (issuing the 16-byte stores from the epilogue threads cost the kernel 19 %).  Two warps on two different
    // schedulers share the blocks; the staging buffers are released (sempty) by the MMA commit AND by both.
    if (aops) {
      const size_t ps = tt_aops_part_bytes(B, G);
      const int64_t nT4 = tt_nT4(G);
      const uint8_t* stg = smem + L::STG;
      for (int j = 0; j < ntiles; ++j) {
        tc::mbar_wait(sfull, j & 1);
        uint8_t* base = aops + (((size_t)blockIdx.x * 8) * nT4 + (t0 + j)) * 4 * 256;
#pragma unroll
        for (int k = warp - 2; k < 2 * L::NSTG; k += 2) {  // warp 2: the hi blocks, warp 3: the lo blocks
          const int part = k;  // 0/1 = g1 s1 hi/lo, 2/3 = p 2^P hi/lo, 4/5 = -dpi 2^D hi/lo
          uint4 v[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] = *reinterpret_cast<const uint4*>(stg + k * TT_ABLK + (lane + 32 * i) * 16);
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const int c = lane + 32 * i, cgc = c >> 7, rr = c & 127;  // chunk c of the block = (column group, row)
            // streaming store (evict-first): 264 MB per launch must not push the weight tiles and h out of L2
            __stcs(reinterpret_cast<uint4*>(base + part * ps + ((size_t)(rr >> 4) * nT4 * 4 + cgc) * 256 + (rr & 15) * 16), v[i]);
          }
        }
        __syncwarp();
        if (lane == 0) tc::mbar_arrive(sempty);
      }
    } else {
      for (int j = 0; j < ntiles; ++j) {
        tc::mbar_wait(sfull, j & 1);
        __syncwarp();
        if (lane == 0) tc::mbar_arrive(sempty);
      }
    }
  } else if (warp >= 4) {
    // ===== epilogue: thread = (row r = 32*(warp%4) + lane, column group cg of 8 genes)
    const int q = warp & 3, cg = (warp - 4) >> 2;
    const int r = 32 * q + lane;
    const bool rvalid = r < valid;
    const int64_t grow = rvalid ? (row_idx ? row_idx[r0 + r] : r0 + r) : 0;
    const void* xrow = XK == XK_U8 ? (const void*)(reinterpret_cast<const uint8_t*>(x) + grow * ldx)
                                   : (const void*)(reinterpret_cast<const float*>(x) + grow * ldx);
    float krow = 0.f, elib_inv = 0.f, s1 = 1.f, inv_s1 = 1.f;
    if (rvalid) {
      float m = -INFINITY, l = 0.f;
      for (int sp = 0; sp < nsplit_lse; ++sp) {
        const float2 ml = *reinterpret_cast<const float2*>(lse_part + ((int64_t)sp * B + r0 + r) * 2);
        if (ml.x == -INFINITY) continue;
        const float mn = fmaxf(m, ml.x);
        l = (m == -INFINITY ? 0.f : l * expf(m - mn)) + ml.y * expf(ml.x - mn);
        m = mn;
      }
      const float lib = library[r0 + r];
      krow = (lib - m) * fm::LOG2E - log2f(l);
      elib_inv = expf(-lib);
      // |g1| <= exp(library) (sum of the row's counts): scale so the bound maps to 2^14
      int e = (int)ceilf(lib * fm::LOG2E);
      e = max(-100, min(100, e));
      s1 = tc::pow2i(14 - e);
      inv_s1 = tc::pow2i(e - 14);
    }
    // staged-operand factors: softmax p = mu exp(-library) x 2^14; -dpi x 2^10
    const float kP = elib_inv * tc::pow2i(TC_P_SCALE_LOG2), kD = -tc::pow2i(TC_DP_SCALE_LOG2);
    const uint32_t lane_sel = (uint32_t)(32 * q) << 16;
    uint8_t* stg = smem + L::STG;
    float nll = 0.f, S = 0.f;
    // counts of this thread's 8 genes of a tile.  The loads for tile j + 1 are issued at the top of tile j: they
    // land during tile j's math, so the fence.proxy.async at the end of the tile (which drains this thread's
    // outstanding loads) never waits on them.
    // counts of this thread's 8 genes of a tile.  The loads for tile j + 1 are issued at the top of tile j: they
    // land during tile j's math, so the fence.proxy.async at the end of the tile (which drains this thread's
    // outstanding loads) never waits on them.  The first use of tile j's counts right behind the issue of tile
    // j + 1's loads does wait for those too (one scoreboard): ~1200 cycles per tile in the clock64 trace
    // (SCVI_K2_TRACE, tools/k2_trace.py).  Consuming the counts first removes that wait but not a cycle of the
    // tile period (6400 cycles either way: the math phase stretches from ~3000 to ~3800 cycles, r02f trace) --
    // the warps of a scheduler then run their dependent MUFU chains in phase instead of staggered.  Also measured
    // and not kept: reading tile j + 1's logits out of TMEM before reporting tile j's staged operands (0.487 ms
    // per step against 0.464).
    XRaw<XK> xcur, xnext;
    if (ntiles > 0) tt_load_x8<XK>(xrow, (int64_t)t0 * TC_NT + 8 * cg, G, rvalid, xvec, xcur);
    for (int j = 0; j < ntiles; ++j) {
      const int s = j & 1;
      const int trw = warp == 4 ? 0 : warp == 19 ? 1 : 3;
      if (trw < 3) K2_TR(trw, j, 6);
      if (j + 1 < ntiles) tt_load_x8<XK>(xrow, (int64_t)(t0 + j + 1) * TC_NT + 8 * cg, G, rvalid, xvec, xnext);
      if (trw < 3) K2_TR(trw, j, 7);
      float xq[8];
      int xi[8];
      const bool lean = tt_prep_x8<XK>(xcur, xq, xi);
#ifdef SCVI_K2_TRACE
      asm volatile("" ::"f"(xq[0]), "f"(xq[7]) : "memory");
#endif
      if (trw < 3) K2_TR(trw, j, 0);
      tc::mbar_wait(&tfull[s], (j >> 1) & 1);
      tc::mbar_wait(&wfull[s], (j >> 1) & 1);  // acquire the bulk-copied gene constants
      if (trw < 3) K2_TR(trw, j, 1);
      tc::fence_after_sync();
      float sv[8], pv[8];
      tc::tmem_ld8(tmem_base + lane_sel + (s * 2 + 0) * TC_NT + 8 * cg, sv);
      if (ZINB) tc::tmem_ld8(tmem_base + lane_sel + (s * 2 + 1) * TC_NT + 8 * cg, pv);
      tc::tmem_ld_wait();
      if (trw < 3) K2_TR(trw, j, 2);
      tc::fence_before_sync();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(&tempty[s]);  // logits buffer drained
      const float* gc = reinterpret_cast<const float*>(smem + L::W0 + s * L::WSTAGE + NH * 2 * TC_BLK32) + 8 * cg * TC_GC;
      const float* tab = reinterpret_cast<const float*>(smem + L::W0 + s * L::WSTAGE + NH * 2 * TC_BLK32 + TC_GC32_BYTES) +
                         8 * cg * fm::LGT;
      float a1[8], a2[8], a3[8];
      if (XK != XK_F32 && lean)
        tt_k2_elems<ZINB, true>(xq, xi, sv, pv, gc, tab, s_lgx1, krow, eps, s1, kP, kD, nll, S, a1, a2, a3);
      else
        tt_k2_elems<ZINB, false>(xq, xi, sv, pv, gc, tab, s_lgx1, krow, eps, s1, kP, kD, nll, S, a1, a2, a3);
      uint4 hi1, lo1, hi2, lo2, hi3, lo3;
      tc::split8(a1, hi1, lo1);
      tc::split8(a2, hi2, lo2);
      if (ZINB) tc::split8(a3, hi3, lo3);
#ifdef SCVI_K2_TRACE
      asm volatile("" ::"r"(hi1.x), "r"(lo2.w), "r"(hi3.y) : "memory");  // the math is done before the stamp
#endif
      if (trw < 3) K2_TR(trw, j, 3);
      tc::mbar_wait(sempty, (j & 1) ^ 1);  // output MMAs of the previous tile have read the staging buffers
      if (trw < 3) K2_TR(trw, j, 4);
      const uint32_t off = cg * (TC_M * 16) + r * 16;
      *reinterpret_cast<uint4*>(stg + 0 * TT_ABLK + off) = hi1;
      *reinterpret_cast<uint4*>(stg + 1 * TT_ABLK + off) = lo1;
      *reinterpret_cast<uint4*>(stg + 2 * TT_ABLK + off) = hi2;
      *reinterpret_cast<uint4*>(stg + 3 * TT_ABLK + off) = lo2;
      if (ZINB) {
        *reinterpret_cast<uint4*>(stg + 4 * TT_ABLK + off) = hi3;
        *reinterpret_cast<uint4*>(stg + 5 * TT_ABLK + off) = lo3;
      }
      tc::fence_proxy_async_smem();
      if (trw < 3) K2_TR(trw, j, 5);
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(sfull);
      xcur = xnext;
    }
    // per-row partials of this gene split: combine the 4 column groups
    s_red[(cg * TC_M + r) * 2 + 0] = nll;
    s_red[(cg * TC_M + r) * 2 + 1] = S;
    asm volatile("bar.sync 1, %0;" ::"n"(32 * TT_EPI_WARPS) : "memory");
    if (cg == 0 && rvalid) {
      float2 o = make_float2(0.f, 0.f);
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        o.x += s_red[(c * TC_M + r) * 2 + 0];
        o.y += s_red[(c * TC_M + r) * 2 + 1];
      }
      *reinterpret_cast<float2*>(out_part + ((int64_t)blockIdx.y * B + r0 + r) * 2) = o;
    }
    // dh partials: thread (r, cg) owns hidden columns 32*cg .. 32*cg+31
    if (ntiles > 0) {
      tc::mbar_wait(accfull, 0);
      tc::fence_after_sync();
    }
    float* da = dh_part + (((int64_t)blockIdx.y * 2 + 0) * B + r0 + r) * TC_H + 32 * cg;
    float* db = dh_part + (((int64_t)blockIdx.y * 2 + 1) * B + r0 + r) * TC_H + 32 * cg;
    const float k1 = inv_s1 * (1.f / TC_SCALE_W), k2 = tc::pow2i(-TC_P_SCALE_LOG2) * (1.f / TC_SCALE_W),
                k3 = tc::pow2i(-TC_DP_SCALE_LOG2) * (1.f / TC_SCALE_W);
#pragma unroll 1
    for (int c = 0; c < 4; ++c) {
      float v1[8], v2[8], v3[8];
      if (ntiles > 0) {
        tc::tmem_ld8(ACC1 + lane_sel + 32 * cg + 8 * c, v1);
        tc::tmem_ld8(ACC2 + lane_sel + 32 * cg + 8 * c, v2);
        if (ZINB) tc::tmem_ld8(ACC3 + lane_sel + 32 * cg + 8 * c, v3);
        tc::tmem_ld_wait();
      } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) v1[i] = v2[i] = v3[i] = 0.f;
      }
      if (rvalid) {
        float oa[8], ob[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          oa[i] = v1[i] * k1 + (ZINB ? v3[i] * k3 : 0.f);
          ob[i] = v2[i] * k2;
        }
        *reinterpret_cast<float4*>(da + 8 * c) = make_float4(oa[0], oa[1], oa[2], oa[3]);
        *reinterpret_cast<float4*>(da + 8 * c + 4) = make_float4(oa[4], oa[5], oa[6], oa[7]);
        *reinterpret_cast<float4*>(db + 8 * c) = make_float4(ob[0], ob[1], ob[2], ob[3]);
        *reinterpret_cast<float4*>(db + 8 * c + 4) = make_float4(ob[4], ob[5], ob[6], ob[7]);
      }
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 2) tc::tmem_dealloc(tmem_base, 512);
}

// reconst[b] = sum_split nll; row_stats[b] = (max, sumexp, exp(library)/sumexp, S);
// dh[b, :] = u_b * (sum_split dh_a - S * sum_split dh_b)        (u_b = g_reconst[b], or 1)
__global__ void tc_train_combine_kernel(const float* __restrict__ lse_part, int nsplit_lse,
                                        const float* __restrict__ out_part, const float* __restrict__ dh_part,
                                        int nsplit, const float* __restrict__ library,
                                        const float* __restrict__ g_reconst, int64_t B, float* __restrict__ reconst,
                                        float* __restrict__ row_stats, float* __restrict__ dh, int64_t lddh,
                                        float* __restrict__ fmax_part /*[blocks][3], nullable*/) {
  // 32 threads per row (one float4 of dh each), 8 rows per block
  __shared__ float s_f[8][3];
  const int64_t b = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (fmax_part) {  // block maxima of the per-row factors (u 2^(e-14), u |S| 2^-P, u): operand scales of the dW GEMM
    if (lane < 3) s_f[threadIdx.x >> 5][lane] = 0.f;
    __syncthreads();
  }
  if (b >= B) {
    if (fmax_part) {
      __syncthreads();
      if (threadIdx.x < 3) {
        float m = 0.f;
        for (int r = 0; r < 8; ++r) m = fmaxf(m, s_f[r][threadIdx.x]);
        fmax_part[(int64_t)blockIdx.x * 3 + threadIdx.x] = m;
      }
    }
    return;
  }
  float m = -INFINITY, l = 0.f;
  for (int sp = 0; sp < nsplit_lse; ++sp) {
    const float2 ml = *reinterpret_cast<const float2*>(lse_part + ((int64_t)sp * B + b) * 2);
    if (ml.x == -INFINITY) continue;
    const float mn = fmaxf(m, ml.x);
    l = (m == -INFINITY ? 0.f : l * expf(m - mn)) + ml.y * expf(ml.x - mn);
    m = mn;
  }
  float nll = 0.f, S = 0.f;
#pragma unroll 8
  for (int sp = 0; sp < nsplit; ++sp) {
    const float2 o = *reinterpret_cast<const float2*>(out_part + ((int64_t)sp * B + b) * 2);
    nll += o.x;
    S += o.y;
  }
  if (lane == 0) {
    reconst[b] = nll;
    *reinterpret_cast<float4*>(row_stats + b * 4) = make_float4(m, l, expf(library[b]) / l, S);
    if (fmax_part) {
      const float u = fabsf(g_reconst ? g_reconst[b] : 1.f);
      const int e = max(-100, min(100, (int)ceilf(library[b] * fm::LOG2E)));  // as the fused kernel's operand scale
      s_f[threadIdx.x >> 5][0] = u * tc::pow2i(e - 14);
      s_f[threadIdx.x >> 5][1] = u * fabsf(S) * tc::pow2i(-TC_P_SCALE_LOG2);
      s_f[threadIdx.x >> 5][2] = u;
    }
  }
  if (dh) {
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f), c = a;
#pragma unroll 8
    for (int sp = 0; sp < nsplit; ++sp) {
      const float4 va = *reinterpret_cast<const float4*>(dh_part + (((int64_t)sp * 2 + 0) * B + b) * TC_H + 4 * lane);
      const float4 vb = *reinterpret_cast<const float4*>(dh_part + (((int64_t)sp * 2 + 1) * B + b) * TC_H + 4 * lane);
      a.x += va.x; a.y += va.y; a.z += va.z; a.w += va.w;
      c.x += vb.x; c.y += vb.y; c.z += vb.z; c.w += vb.w;
    }
    const float u = g_reconst ? g_reconst[b] : 1.f;
    *reinterpret_cast<float4*>(dh + b * lddh + 4 * lane) =
        make_float4(u * (a.x - S * c.x), u * (a.y - S * c.y), u * (a.z - S * c.z), u * (a.w - S * c.w));
  }
  if (fmax_part) {
    __syncthreads();
    if (threadIdx.x < 3) {
      float mx = 0.f;
      for (int r = 0; r < 8; ++r) mx = fmaxf(mx, s_f[r][threadIdx.x]);
      fmax_part[(int64_t)blockIdx.x * 3 + threadIdx.x] = mx;
    }
  }
}

// ---------------------------------------------------------------------------------------
// host: K2 launcher (called by tc_decoder_nll_fwd in decoder_tc.cu when dh is requested)
// ---------------------------------------------------------------------------------------
template <bool ZINB, int XK, bool AOPS>
static int tt_launch_k2(const float* h, int64_t ldh, int64_t B, int64_t G, const uint8_t* packed32, const void* x,
                        int64_t ldx, const int64_t* row_idx, const float* library, const float* lse, int nsplit_lse,
                        int nsplit, int tiles_per_split, int xvec, float* out_part, float* dh_part, uint8_t* aops,
                        cudaStream_t st) {
  auto k = tc_nll_dh_kernel<ZINB, XK, AOPS>;
  constexpr int smem = (int)K2Smem<ZINB, AOPS>::TOTAL + 128;
  static std::atomic<uint64_t> optin{0};
  if (int rc = smem_optin((const void*)k, smem, optin, "tc_nll_dh")) return rc;
  dim3 grid((unsigned)cdiv(B, TC_M), (unsigned)nsplit);
  k<<<grid, TT_THREADS, smem, st>>>(h, ldh, B, G, packed32, x, ldx, row_idx, library, lse, nsplit_lse, tiles_per_split,
                                    1e-8f, xvec, out_part, dh_part, aops);
  return SCVI_OK;
}

// xkind: XK_F32 / XK_F32_INT / XK_U8 (the count operand; ldx in elements of that type)
int tc_launch_nll_dh(bool zinb, int xkind, const float* h, int64_t ldh, int64_t B, int64_t G, const uint8_t* packed32,
                     const void* x, int64_t ldx, const int64_t* row_idx, const float* library, const float* lse,
                     int nsplit_lse, int nsplit, int tiles_per_split, float* out_part, float* dh_part,
                     const float* g_reconst, float* reconst, float* row_stats, float* dh, int64_t lddh,
                     uint8_t* aops, float* fmax_part, cudaStream_t st) {
  const int xvec = xkind == XK_U8 ? ((ldx % 8 == 0) && ((uintptr_t)x % 8 == 0))
                                  : ((ldx % 4 == 0) && (G % 4 == 0) && ((uintptr_t)x % 16 == 0));
  int rc;
#define SCVI_K2(Z, K, A) \
  tt_launch_k2<Z, K, A>(h, ldh, B, G, packed32, x, ldx, row_idx, library, lse, nsplit_lse, nsplit, tiles_per_split, xvec, out_part, dh_part, aops, st)
  const bool keep = aops != nullptr && xkind != XK_F32;
  if (zinb) rc = xkind == XK_U8 ? (keep ? SCVI_K2(true, XK_U8, true) : SCVI_K2(true, XK_U8, false))
                 : xkind == XK_F32_INT ? (keep ? SCVI_K2(true, XK_F32_INT, true) : SCVI_K2(true, XK_F32_INT, false))
                                       : SCVI_K2(true, XK_F32, false);
  else rc = xkind == XK_U8 ? (keep ? SCVI_K2(false, XK_U8, true) : SCVI_K2(false, XK_U8, false))
            : xkind == XK_F32_INT ? (keep ? SCVI_K2(false, XK_F32_INT, true) : SCVI_K2(false, XK_F32_INT, false))
                                  : SCVI_K2(false, XK_F32, false);
#undef SCVI_K2
  if (rc) return rc;
  SCVI_LAUNCH_CHECK("scvi_decoder_nll_fwd(tc nll+dh)");
  tc_train_combine_kernel<<<(unsigned)cdiv(B, 8), 256, 0, st>>>(lse, nsplit_lse, out_part, dh_part, nsplit, library,
                                                               g_reconst, B, reconst, row_stats, dh, lddh, fmax_part);
  SCVI_LAUNCH_CHECK("scvi_decoder_nll_fwd(tc combine)");
  return SCVI_OK;
}

}  // namespace scvi

// =======================================================================================
// K3: dWs, dWp, dbs, dbp, dpx_r  (genes-owned, transposed tiles: TMEM lane = gene, column = row)
// =======================================================================================
namespace scvi {

// h fp32 [B, H] -> per 32-row tile {h hi block, h lo block, row constants}; also per-tile bounds that fix
// the staged-operand scales of K3 (which takes their maximum -- no atomics, no zero-fill):
// scal[t][0] = max_b 3 u_b exp(library_b) >= |ds|, scal[t][1] = max_b |u_b| over the rows b of tile t.
__global__ void __launch_bounds__(512)
    tc_pack_h32_kernel(const float* __restrict__ h, int64_t ldh, int64_t B, const float* __restrict__ row_stats,
                       const float* __restrict__ g_reconst, const int64_t* __restrict__ row_idx,
                       uint8_t* __restrict__ hpack, float* __restrict__ scal /*[nR][2]*/) {
  const int64_t t = blockIdx.x;
  const int r = threadIdx.x & 31, kc = threadIdx.x >> 5;
  const int64_t b = t * TC_NT + r;
  uint8_t* base = hpack + t * TC_H32_STAGE_BYTES;
  float v[8];
  if (b < B) {
    const float4 a = *reinterpret_cast<const float4*>(h + b * ldh + 8 * kc);
    const float4 c = *reinterpret_cast<const float4*>(h + b * ldh + 8 * kc + 4);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = c.x; v[5] = c.y; v[6] = c.z; v[7] = c.w;
#pragma unroll
    for (int q = 0; q < 8; ++q) v[q] = fminf(fmaxf(v[q] * TC_SCALE_H, -60000.f), 60000.f);
  } else {
#pragma unroll
    for (int q = 0; q < 8; ++q) v[q] = 0.f;
  }
  uint4 hi, lo;
  tc::split8(v, hi, lo);
  *reinterpret_cast<uint4*>(base + kc * TC_NT * 16 + r * 16) = hi;
  *reinterpret_cast<uint4*>(base + TC_BLK32 + kc * TC_NT * 16 + r * 16) = lo;
  if (kc == 0) {
    float* rc = reinterpret_cast<float*>(base + 2 * TC_BLK32) + r * TC_ROWC;
    float krow = 0.f, u = 0.f, uSe = 0.f, ok = 0.f, b_ds = 0.f, b_dp = 0.f;
    int64_t rid = 0;
    if (b < B) {
      const float4 st = *reinterpret_cast<const float4*>(row_stats + b * 4);  // (max, sumexp, exp(lib)/sumexp, S)
      u = g_reconst ? g_reconst[b] : 1.f;
      krow = -st.x * fm::LOG2E + log2f(st.z);
      const float L = st.z * st.y;  // exp(library)
      uSe = u * st.w / L;
      ok = 1.f;
      rid = row_idx ? row_idx[b] : b;
      b_ds = 3.f * fabsf(u) * L;
      b_dp = fabsf(u);
    }
    b_ds = warp_max(b_ds);  // kc == 0 is exactly warp 0
    b_dp = warp_max(b_dp);
    if (r == 0) *reinterpret_cast<float2*>(scal + 2 * t) = make_float2(b_ds, b_dp);
    rc[0] = krow; rc[1] = u; rc[2] = uSe; rc[3] = ok;
    rc[4] = __uint_as_float((uint32_t)(rid & 0xffffffffu));
    rc[5] = __uint_as_float((uint32_t)((uint64_t)rid >> 32));
    rc[6] = 0.f; rc[7] = 0.f;
  }
}

template <bool ZINB>
struct K3Smem {
  static constexpr int NH = ZINB ? 2 : 1;
  static constexpr uint32_t W = 0;                              // [head][hi, lo] 32 KB blocks
  static constexpr uint32_t HST = NH * 2 * TC_OPER_BYTES;       // 2 stages of {h hi, h lo, row constants}
  static constexpr uint32_t STG = HST + 2 * TC_H32_STAGE_BYTES; // [head][hi, lo] staged ds^T / dp^T
  static constexpr uint32_t BARS = STG + NH * 2 * TT_ABLK;
  static constexpr uint32_t TMEM_PTR = BARS + 16 * 8;
  static constexpr uint32_t RED = TMEM_PTR + 16;                // [3][4 row groups][128 genes]
  static constexpr uint32_t DGTAB = RED + 3 * 4 * TC_M * 4;     // [128 genes][DGT_LD] digamma-difference table
  static constexpr uint32_t DGTAB_BYTES = TC_M * fm::DGT_LD * 4;
  static constexpr uint32_t TOTAL = DGTAB + DGTAB_BYTES;
};
static_assert(K3Smem<true>::DGTAB % 16 == 0 && K3Smem<true>::DGTAB_BYTES % 16 == 0, "bulk copy alignment");
static_assert(K3Smem<true>::TOTAL + 128 <= 232448, "K3 shared memory");

// the 8 elements (8 rows x this thread's gene) of one tile: derivatives, column sums, staged operands.
// Padded rows carry u = 0 (every product below is an exact zero), padded genes are never stored: no masks.
template <bool ZINB, bool LEAN>
__device__ __forceinline__ void tt_k3_elems(const float (&xq)[8], const int (&xi)[8], const float (&sv)[8],
                                            const float (&pv)[8], const float* __restrict__ rowc,
                                            const float* __restrict__ dgt, const fm::GeneK& gk, float bsv, float bpv,
                                            float eps, float sc_ds, float sc_dp, float& cs, float& cp, float& cr,
                                            float (&a1)[8], float (&a2)[8]) {
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const float4 rc = *reinterpret_cast<const float4*>(rowc + i * TC_ROWC);  // krow, u, u*S*exp(-lib), valid
    const float sl = fmaf(sv[i], TC_RESCALE, bsv);
    const float pl = ZINB ? fmaf(pv[i], TC_RESCALE, bpv) : 0.f;
    float g1, dpi, dth, mu;
    if (LEAN)
      fm::elem_bwd_lean<ZINB>(xq[i], dgt[xi[i]], sl, pl, rc.x, gk, eps, g1, dpi, dth, mu);
    else
      fm::elem_bwd_tab<ZINB>(xq[i], sl, pl, rc.x, gk, eps, dgt, g1, dpi, dth, mu);
    const float ds = fmaf(rc.y, g1, -mu * rc.z);
    const float dp = -rc.y * dpi;
    cs += ds;
    cp += dp;
    cr = fmaf(-rc.y * gk.theta, dth, cr);
    a1[i] = ds * sc_ds;
    a2[i] = dp * sc_dp;
  }
}

template <bool ZINB, int XK>
__global__ void __launch_bounds__(TT_THREADS, 1)
    tc_dw_kernel(int64_t B, int64_t G, const uint8_t* __restrict__ packed128, const float* __restrict__ gconst,
                 const uint8_t* __restrict__ hpack, const float* __restrict__ scal, const void* __restrict__ x,
                 int64_t ldx, int tiles_per_split, float eps, float* __restrict__ slabW /*[splits][NH][G][H]*/,
                 float* __restrict__ slabv /*[splits][3][G]*/) {
  using L = K3Smem<ZINB>;
  constexpr int NH = L::NH;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((128u - (tc::smem_u32(smem_raw) & 127u)) & 127u);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::BARS);
  uint64_t *hfull = bars, *hempty = bars + 2, *tfull = bars + 4, *tempty = bars + 6;
  uint64_t *sfull = bars + 8, *sempty = bars + 9, *accfull = bars + 10, *wfull = bars + 11;
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(smem + L::TMEM_PTR);
  float* s_red = reinterpret_cast<float*>(smem + L::RED);
  float* s_bound = reinterpret_cast<float*>(smem + L::TMEM_PTR + 8);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int64_t g0 = (int64_t)blockIdx.x * TC_M;
  const int nR = (int)((B + TC_NT - 1) / TC_NT);
  const int rt0 = blockIdx.y * tiles_per_split;
  const int rt1 = min(nR, rt0 + tiles_per_split);
  const int ntiles = max(0, rt1 - rt0);

  if (warp == 1 && lane == 0) {
    for (int i = 0; i < 2; ++i) {
      tc::mbar_init(&hfull[i], 1);
      tc::mbar_init(&hempty[i], 1);
      tc::mbar_init(&tfull[i], 1);
      tc::mbar_init(&tempty[i], TT_EPI_WARPS);
    }
    tc::mbar_init(sfull, TT_EPI_WARPS);
    tc::mbar_init(sempty, 1);
    tc::mbar_init(accfull, 1);
    tc::mbar_init(wfull, 1);
    tc::fence_mbar_init();
  }
  if (warp == 2) tc::tmem_alloc(s_tmem, 512);
  if (warp == 3) {  // global operand bounds = max over the per-tile bounds
    float b0 = 0.f, b1 = 0.f;
    for (int i = lane; i < nR; i += 32) {
      const float2 v = *reinterpret_cast<const float2*>(scal + 2 * i);
      b0 = fmaxf(b0, v.x);
      b1 = fmaxf(b1, v.y);
    }
    b0 = warp_max(b0);
    b1 = warp_max(b1);
    if (lane == 0) {
      s_bound[0] = b0;
      s_bound[1] = b1;
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem_base = *s_tmem;
  // TMEM columns: logits^T [buf][head] 32 each at 0..127; dWs accumulator at 128, dWp at 256

  if (warp == 0) {
    if (lane == 0 && ntiles > 0) {
      // resident weight tile of this CTA's 128 genes: NH x {hi, lo} blocks, contiguous in packed128
      tc::mbar_arrive_expect_tx(wfull, NH * 2 * TC_OPER_BYTES + L::DGTAB_BYTES);
      const uint8_t* src = packed128 + (size_t)blockIdx.x * NH * 2 * TC_OPER_BYTES;
      for (int i = 0; i < NH * 2; ++i)
        tc::bulk_g2s(smem + L::W + i * TC_OPER_BYTES, src + (size_t)i * TC_OPER_BYTES, TC_OPER_BYTES, wfull);
      // this CTA's 128 rows of the dgd table (stored behind the gene constants and the bias copy)
      const int nT128 = (int)((G + TC_M - 1) / TC_M);
      tc::bulk_g2s(smem + L::DGTAB, gconst + (size_t)nT128 * TC_M * (TC_GC + 1) + (size_t)blockIdx.x * TC_M * fm::DGT_LD,
                   L::DGTAB_BYTES, wfull);
      for (int j = 0; j < ntiles; ++j) {
        const int s = j & 1;
        tc::mbar_wait(&hempty[s], ((j >> 1) & 1) ^ 1);
        tc::mbar_arrive_expect_tx(&hfull[s], TC_H32_STAGE_BYTES);
        tc::bulk_g2s(smem + L::HST + s * TC_H32_STAGE_BYTES, hpack + (size_t)(rt0 + j) * TC_H32_STAGE_BYTES,
                     TC_H32_STAGE_BYTES, &hfull[s]);
      }
    }
  } else if (warp == 1) {
    if (ntiles > 0) {  // warp-uniform issue loop (tc.cuh)
      constexpr uint32_t idescL = tc::idesc_f16_f32(TC_M, TC_NT, 0, 0);  // logits^T: 128 genes x 32 rows
      constexpr uint32_t idescO = tc::idesc_f16_f32(TC_M, TC_H, 0, 1);   // dW: 128 genes x 128 hidden, B MN-major
      const uint32_t leader = tc::elect_one();
      const uint32_t wbase = tc::smem_u32(smem + L::W);
      const uint32_t stg = tc::smem_u32(smem + L::STG);
      const uint32_t hst = tc::smem_u32(smem + L::HST);
      // A of the logits^T GEMM = resident weight tile of head hd, K-major (128 genes x 128 hidden)
      // A of the dW GEMM = staged ds^T / dp^T (K-major: 128 genes x 32 rows)
      tc::Oper wA[NH], sA[NH];
#pragma unroll
      for (int hd = 0; hd < NH; ++hd) {
        wA[hd] = tc::make_oper(wbase + hd * 2 * TC_OPER_BYTES, wbase + hd * 2 * TC_OPER_BYTES + TC_OPER_BYTES, TC_M * 16, 128,
                               2 * TC_M * 16);
        sA[hd] = tc::make_oper(stg + (hd * 2) * TT_ABLK, stg + (hd * 2 + 1) * TT_ABLK, TC_M * 16, 128, 2 * TC_M * 16);
      }
      // h tile of stage 0 (stage 1 = + TC_H32_STAGE_BYTES): K-major B of the logits^T GEMM (N = 32 rows, K = hidden)
      // and MN-major B of the dW GEMM (N = hidden, K = rows)
      const tc::Oper hL = tc::make_oper(hst, hst + TC_BLK32, TC_NT * 16, 128, 2 * TC_NT * 16);
      const tc::Oper hO = tc::make_oper(hst, hst + TC_BLK32, 128, TC_NT * 16, 256);
      auto stage_of = [](const tc::Oper& o, int s) {
        tc::Oper r = o;
        r.hi_lo += s * (TC_H32_STAGE_BYTES >> 4);
        r.lo_lo += s * (TC_H32_STAGE_BYTES >> 4);
        return r;
      };
      tc::mbar_wait(wfull, 0);
      auto issue_logits = [&](int j) {
        const int s = j & 1;
        tc::mbar_wait(&hfull[s], (j >> 1) & 1);
        tc::mbar_wait(&tempty[s], ((j >> 1) & 1) ^ 1);
        tc::fence_after_sync();
        if (leader) {
#pragma unroll
          for (int hd = 0; hd < NH; ++hd)
            tc::mma3<TC_H / 16>(tmem_base + (s * 2 + hd) * TC_NT, wA[hd], stage_of(hL, s), idescL, false);
          tc::mma_commit(&tfull[s]);
        }
        __syncwarp();
      };
      issue_logits(0);
      for (int j = 0; j < ntiles; ++j) {
        if (j + 1 < ntiles) issue_logits(j + 1);
        const int s = j & 1;
        tc::mbar_wait(sfull, j & 1);
        tc::fence_after_sync();
        if (leader) {
#pragma unroll
          for (int hd = 0; hd < NH; ++hd)
            tc::mma3<TC_NT / 16>(tmem_base + 128 + hd * TC_H, sA[hd], stage_of(hO, s), idescO, j > 0);
          tc::mma_commit(sempty);
          tc::mma_commit(&hempty[s]);
        }
        __syncwarp();
      }
      if (leader) tc::mma_commit(accfull);
      __syncwarp();
    }
  } else if (warp >= 4) {
    // ===== epilogue: thread = (gene lane gl, row group rg of 8 rows)
    const int q = warp & 3, rg = (warp - 4) >> 2;
    const int gl = 32 * q + lane;
    const int64_t g = g0 + gl;
    const bool gvalid = g < G;
    const float* gcp = gconst + g * TC_GC;  // the table is padded to whole 128-gene tiles
    const fm::GeneK gk = load_gene_k(gcp);
    const float bsv = gcp[3], bpv = gcp[4];
    // staged-operand scales from the global bounds (powers of two)
    const float bound_ds = fmaxf(s_bound[0], 1e-30f), bound_dp = fmaxf(s_bound[1], 1e-30f);
    const int e_ds = max(-100, min(100, (int)ceilf(log2f(bound_ds)))), e_dp = max(-100, min(100, (int)ceilf(log2f(bound_dp))));
    const float sc_ds = tc::pow2i(14 - e_ds), sc_dp = tc::pow2i(TC_DP_SCALE_LOG2 - e_dp);
    const uint32_t lane_sel = (uint32_t)(32 * q) << 16;
    uint8_t* stg = smem + L::STG;
    const float* dgt = reinterpret_cast<const float*>(smem + L::DGTAB) + gl * fm::DGT_LD;
    if (ntiles > 0) tc::mbar_wait(wfull, 0);  // acquire the bulk-copied dgd table (it lands with the weight tile)
    float cs = 0.f, cp = 0.f, cr = 0.f;
    for (int j = 0; j < ntiles; ++j) {
      const int s = j & 1;
      tc::mbar_wait(&hfull[s], (j >> 1) & 1);  // acquire the row constants of this tile
      const float* rowc = reinterpret_cast<const float*>(smem + L::HST + s * TC_H32_STAGE_BYTES + 2 * TC_BLK32) + 8 * rg * TC_ROWC;
      float xq[8];
      int xi[8];
      uint32_t xmax = 0u;
      constexpr float MAGIC = 8388608.f;  // 2^23: integer n < 2^23 -> float bits 0x4B000000 | n
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float2 id = *reinterpret_cast<const float2*>(rowc + i * TC_ROWC + 4);
        const int64_t rid = (int64_t)(((uint64_t)__float_as_uint(id.y) << 32) | (uint64_t)__float_as_uint(id.x));
        // padded rows carry row id 0; padded genes read nothing
        if (XK == XK_U8) {
          const uint32_t b = gvalid ? (uint32_t)__ldg(reinterpret_cast<const uint8_t*>(x) + rid * ldx + g) : 0u;
          xi[i] = (int)b;
          xmax = max(xmax, b);
          xq[i] = __uint_as_float(0x4B000000u | b) - MAGIC;
        } else {
          xq[i] = gvalid ? ld_stream(reinterpret_cast<const float*>(x) + rid * ldx + g) : 0.f;
          xi[i] = 0;
        }
      }
      if (XK == XK_F32_INT) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const uint32_t n = __float_as_uint(xq[i] + MAGIC) ^ 0x4B000000u;  // the count itself when it is an integer < 2^23
          xmax = max(xmax, n);
          xi[i] = (int)(n & 63u);
        }
      }
      const bool lean = XK != XK_F32 && xmax < (uint32_t)fm::DGT;
      tc::mbar_wait(&tfull[s], (j >> 1) & 1);
      tc::fence_after_sync();
      float sv[8], pv[8];
      tc::tmem_ld8(tmem_base + lane_sel + (s * 2 + 0) * TC_NT + 8 * rg, sv);
      if (ZINB) tc::tmem_ld8(tmem_base + lane_sel + (s * 2 + 1) * TC_NT + 8 * rg, pv);
      tc::tmem_ld_wait();
      tc::fence_before_sync();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(&tempty[s]);
      float a1[8], a2[8];
      if (lean)
        tt_k3_elems<ZINB, true>(xq, xi, sv, pv, rowc, dgt, gk, bsv, bpv, eps, sc_ds, sc_dp, cs, cp, cr, a1, a2);
      else
        tt_k3_elems<ZINB, false>(xq, xi, sv, pv, rowc, dgt, gk, bsv, bpv, eps, sc_ds, sc_dp, cs, cp, cr, a1, a2);
      uint4 hi1, lo1, hi2, lo2;
      tc::split8(a1, hi1, lo1);
      if (ZINB) tc::split8(a2, hi2, lo2);
      tc::mbar_wait(sempty, (j & 1) ^ 1);
      const uint32_t off = rg * (TC_M * 16) + gl * 16;
      *reinterpret_cast<uint4*>(stg + 0 * TT_ABLK + off) = hi1;
      *reinterpret_cast<uint4*>(stg + 1 * TT_ABLK + off) = lo1;
      if (ZINB) {
        *reinterpret_cast<uint4*>(stg + 2 * TT_ABLK + off) = hi2;
        *reinterpret_cast<uint4*>(stg + 3 * TT_ABLK + off) = lo2;
      }
      tc::fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(sfull);
    }
    // column sums: combine the 4 row groups of each gene
    s_red[(0 * 4 + rg) * TC_M + gl] = cs;
    s_red[(1 * 4 + rg) * TC_M + gl] = cp;
    s_red[(2 * 4 + rg) * TC_M + gl] = cr;
    asm volatile("bar.sync 1, %0;" ::"n"(32 * TT_EPI_WARPS) : "memory");
    if (rg < 3 && gvalid) {  // row group rg finalises quantity rg (dbs, dbp, dpx_r)
      float sum = 0.f;
#pragma unroll
      for (int c = 0; c < 4; ++c) sum += s_red[(rg * 4 + c) * TC_M + gl];
      slabv[((int64_t)blockIdx.y * 3 + rg) * G + g] = sum;
    }
    if (ntiles > 0) {
      tc::mbar_wait(accfull, 0);
      tc::fence_after_sync();
    }
    const float k1 = 1.f / (sc_ds * TC_SCALE_H), k2 = 1.f / (sc_dp * TC_SCALE_H);
#pragma unroll 1
    for (int hd = 0; hd < NH; ++hd) {
      float* o = slabW + (((int64_t)blockIdx.y * NH + hd) * G + g) * TC_H + 32 * rg;
      const float kk = hd == 0 ? k1 : k2;
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        float v[8];
        if (ntiles > 0) {
          tc::tmem_ld8(tmem_base + lane_sel + 128 + hd * TC_H + 32 * rg + 8 * c, v);
          tc::tmem_ld_wait();
        } else {
#pragma unroll
          for (int i = 0; i < 8; ++i) v[i] = 0.f;
        }
        if (gvalid) {
          *reinterpret_cast<float4*>(o + 8 * c) = make_float4(v[0] * kk, v[1] * kk, v[2] * kk, v[3] * kk);
          *reinterpret_cast<float4*>(o + 8 * c + 4) = make_float4(v[4] * kk, v[5] * kk, v[6] * kk, v[7] * kk);
        }
      }
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 2) tc::tmem_dealloc(tmem_base, 512);
}

// sum the row-split slabs in fixed order (deterministic)
static __global__ void tt_bwd_reduce_kernel(const float* __restrict__ slabW, const float* __restrict__ slabv,
                                            int64_t splits, int nheads, int64_t GH, int64_t G,
                                            float* __restrict__ dWs, float* __restrict__ dWp, float* __restrict__ dbs,
                                            float* __restrict__ dbp, float* __restrict__ dpx_r) {
  const int64_t e = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (e < GH) {
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f), p = s;
#pragma unroll 4
    for (int64_t t = 0; t < splits; ++t) {
      const float4 a = *reinterpret_cast<const float4*>(slabW + (t * nheads + 0) * GH + e);
      s.x += a.x; s.y += a.y; s.z += a.z; s.w += a.w;
      if (nheads == 2) {
        const float4 b = *reinterpret_cast<const float4*>(slabW + (t * nheads + 1) * GH + e);
        p.x += b.x; p.y += b.y; p.z += b.z; p.w += b.w;
      }
    }
    if (dWs) *reinterpret_cast<float4*>(dWs + e) = s;
    if (nheads == 2 && dWp) *reinterpret_cast<float4*>(dWp + e) = p;
  }
  const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (g < G) {
    float a = 0.f, b = 0.f, c = 0.f;
#pragma unroll 4
    for (int64_t t = 0; t < splits; ++t) {
      a += slabv[(t * 3 + 0) * G + g];
      b += slabv[(t * 3 + 1) * G + g];
      c += slabv[(t * 3 + 2) * G + g];
    }
    if (dbs) dbs[g] = a;
    if (nheads == 2 && dbp) dbp[g] = b;
    if (dpx_r) dpx_r[g] = c;
  }
}

__global__ void tc_pack_heads_kernel(const float* __restrict__ Ws, const float* __restrict__ bs,
                                     const float* __restrict__ Wp, const float* __restrict__ bp,
                                     const float* __restrict__ px_r, int64_t G, int nheads, float eps,
                                     __half* __restrict__ packed, uint8_t* __restrict__ packed32,
                                     float* __restrict__ gconst);

void tc_decoder_fwd_ws_offsets(int64_t B, int64_t G, int nheads, size_t* off_packed, size_t* off_gconst);
void tc_decoder_fwd_ws_aops(int64_t B, int64_t G, int nheads, size_t* off_fmax, size_t* off_aops);
int64_t tc_decoder_bwd2_ws_bytes(int64_t B, int64_t G, int likelihood);  // decoder_tc_dw2.cu
int tc_decoder_nll_bwd_dw2(const float* h, int64_t ldh, int64_t B, int64_t G, const float* px_r, const void* x,
                           int64_t ldx, const int64_t* row_idx, const float* library, const float* row_stats,
                           const float* g_reconst, int likelihood, float* dWs, float* dbs, float* dWp, float* dbp,
                           float* dpx_r, void* ws, const uint8_t* aops, const float* fmax_part, int xkind,
                           cudaStream_t st);

struct TtBwdPlan {
  int64_t nT, nR;
  int splits, tiles_per_split;
  size_t off_packed, off_gconst, off_hpack, off_scal, off_slabW, off_slabv, total;
};

static TtBwdPlan tt_bwd_plan(int64_t B, int64_t G, int nheads) {
  TtBwdPlan p;
  p.nT = cdiv(G, TC_M);
  p.nR = cdiv(B, TC_NT);
  int64_t s = std::min<int64_t>(std::max<int64_t>(1, 148 / p.nT), p.nR);
  p.tiles_per_split = (int)cdiv(p.nR, s);
  p.splits = (int)cdiv(p.nR, p.tiles_per_split);
  auto al = [](size_t v) { return (v + 255) / 256 * 256; };
  p.off_packed = 0;
  p.off_gconst = al((size_t)p.nT * nheads * 2 * TC_OPER_BYTES);
  p.off_hpack = al(p.off_gconst + (size_t)p.nT * TC_M * TC_GCONST_FLOATS * sizeof(float));
  p.off_scal = al(p.off_hpack + (size_t)p.nR * TC_H32_STAGE_BYTES);
  p.off_slabW = al(p.off_scal + (size_t)p.nR * 2 * sizeof(float));
  p.off_slabv = al(p.off_slabW + (size_t)p.splits * nheads * G * TC_H * sizeof(float));
  p.total = al(p.off_slabv + (size_t)p.splits * 3 * G * sizeof(float));
  return p;
}

int64_t tc_decoder_bwd_ws_bytes(int64_t B, int64_t G, int likelihood) {
  return std::max<int64_t>((int64_t)tt_bwd_plan(B, G, likelihood == SCVI_LIK_ZINB ? 2 : 1).total,
                           tc_decoder_bwd2_ws_bytes(B, G, likelihood));
}

template <bool ZINB, int XK>
static int tt_launch_k3(const TtBwdPlan& p, int64_t B, int64_t G, const uint8_t* packed, const float* gconst,
                        const uint8_t* hpack, const float* scal, const void* x, int64_t ldx, float* slabW,
                        float* slabv, cudaStream_t st) {
  auto k = tc_dw_kernel<ZINB, XK>;
  constexpr int smem = (int)K3Smem<ZINB>::TOTAL + 128;
  static std::atomic<uint64_t> optin{0};
  if (int rc = smem_optin((const void*)k, smem, optin, "tc_dw")) return rc;
  dim3 grid((unsigned)p.nT, (unsigned)p.splits);
  k<<<grid, TT_THREADS, smem, st>>>(B, G, packed, gconst, hpack, scal, x, ldx, p.tiles_per_split, 1e-8f, slabW, slabv);
  return SCVI_OK;
}

// weight-gradient half of the decoder backward on the tensor cores (H == 128)
int tc_decoder_nll_bwd_dw(const float* h, int64_t ldh, int64_t B, int64_t G, const float* Ws, const float* bs,
                          const float* Wp, const float* bp, const float* px_r, const void* x, int64_t ldx,
                          const int64_t* row_idx, const float* row_stats, const float* g_reconst, int likelihood,
                          float* dWs, float* dbs, float* dWp, float* dbp, float* dpx_r, void* ws, const void* fwd_ws,
                          int xkind, const float* library, cudaStream_t st) {
  const bool zinb = likelihood == SCVI_LIK_ZINB;
  const int nheads = zinb ? 2 : 1;
  if (xkind & 16) {  // SCVI_PREC_AOPS: the matching forward kept its operands in fwd_ws -> pure GEMM
    SCVI_REQUIRE(fwd_ws != nullptr, "scvi_decoder_nll_bwd(SCVI_PREC_AOPS): needs the forward workspace");
    size_t off_fmax, off_aops;
    tc_decoder_fwd_ws_aops(B, G, nheads, &off_fmax, &off_aops);
    const uint8_t* f8 = (const uint8_t*)fwd_ws;
    return tc_decoder_nll_bwd_dw2(h, ldh, B, G, px_r, x, ldx, row_idx, library, row_stats, g_reconst, likelihood, dWs, dbs,
                                  dWp, dbp, dpx_r, ws, f8 + off_aops, (const float*)(f8 + off_fmax), xkind & 15, st);
  }
  xkind &= 15;
  SCVI_REQUIRE(ws != nullptr && ((uintptr_t)ws % 256 == 0), "scvi_decoder_nll_bwd(SCVI_PREC_TC): ws must be 256-byte aligned");
  SCVI_REQUIRE(ldh % 4 == 0 && ((uintptr_t)h % 16 == 0), "scvi_decoder_nll_bwd(SCVI_PREC_TC): h must be 16-byte aligned");
  const TtBwdPlan p = tt_bwd_plan(B, G, nheads);
  uint8_t* w8 = (uint8_t*)ws;
  __half* packed = (__half*)(w8 + p.off_packed);
  float* gconst = (float*)(w8 + p.off_gconst);
  uint8_t* hpack = w8 + p.off_hpack;
  float* scal = (float*)(w8 + p.off_scal);
  float* slabW = (float*)(w8 + p.off_slabW);
  float* slabv = (float*)(w8 + p.off_slabv);
  if (fwd_ws) {  // packed weights + gene constants of the matching forward call
    size_t fwd_off_packed, fwd_off_gconst;
    tc_decoder_fwd_ws_offsets(B, G, nheads, &fwd_off_packed, &fwd_off_gconst);
    packed = (__half*)((uint8_t*)fwd_ws + fwd_off_packed);
    gconst = (float*)((uint8_t*)fwd_ws + fwd_off_gconst);
  } else {
    const int64_t items = std::max<int64_t>(p.nT * nheads * TC_KC * TC_N, p.nT * TC_N * fm::DGT_LD);
    tc_pack_heads_kernel<<<(unsigned)cdiv(items, 256), 256, 0, st>>>(Ws, bs, Wp, bp, px_r, G, nheads, 1e-8f, packed, nullptr, gconst);
    SCVI_LAUNCH_CHECK("scvi_decoder_nll_bwd(tc pack W)");
  }
  tc_pack_h32_kernel<<<(unsigned)p.nR, 512, 0, st>>>(h, ldh, B, row_stats, g_reconst, row_idx, hpack, scal);
  SCVI_LAUNCH_CHECK("scvi_decoder_nll_bwd(tc pack h)");
  int rc;
#define SCVI_K3(Z, K) tt_launch_k3<Z, K>(p, B, G, (const uint8_t*)packed, gconst, hpack, scal, x, ldx, slabW, slabv, st)
  if (zinb) rc = xkind == XK_U8 ? SCVI_K3(true, XK_U8) : xkind == XK_F32_INT ? SCVI_K3(true, XK_F32_INT) : SCVI_K3(true, XK_F32);
  else rc = xkind == XK_U8 ? SCVI_K3(false, XK_U8) : xkind == XK_F32_INT ? SCVI_K3(false, XK_F32_INT) : SCVI_K3(false, XK_F32);
#undef SCVI_K3
  if (rc) return rc;
  SCVI_LAUNCH_CHECK("scvi_decoder_nll_bwd(tc dW)");
  const int64_t GH = G * TC_H;
  tt_bwd_reduce_kernel<<<(unsigned)cdiv(std::max<int64_t>(GH / 4, G), 256), 256, 0, st>>>(slabW, slabv, p.splits, nheads, GH, G, dWs, dWp, dbs, dbp, dpx_r);
  SCVI_LAUNCH_CHECK("scvi_decoder_nll_bwd(tc reduce)");
  return SCVI_OK;
}

}  // namespace scvi

#ifdef SCVI_K2_TRACE
extern "C" int scvi_debug_k2_trace(long long* host_out, int n) {
  return (int)cudaMemcpyFromSymbol(host_out, scvi::g_k2_trace, sizeof(long long) * (size_t)n);
}
#endif
