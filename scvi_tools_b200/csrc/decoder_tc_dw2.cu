// Weight-gradient half of the decoder backward as a PURE tensor-core GEMM (SCVI_PREC_AOPS).
//
// The recompute kernel (tc_dw_kernel, decoder_tc_train.cu) re-derives every per-element derivative the fused
// likelihood kernel K2 already had in registers: ~100 us of SFU-bound work per step.  With SCVI_PREC_AOPS K2 keeps its
// staged operands (fp16 hi/lo) in the forward workspace, laid out so that the block a GEMM CTA needs is contiguous,
// and the gradients become
//     dWs = (g1 s1)^T (u/s1 . h)  +  (p 2^P)^T (-u S 2^-P . h)          dbs = the same products with a ones column
//     dWp = (-dpi 2^D)^T (u 2^-D . h)                                    dbp likewise
//     dpx_r = -theta sum_b u_b d(ll)/d(theta)_bg
// (u = loss weight of the row, S = the row's sum of g1, s1 / 2^P / 2^D = K2's operand scales; the row factors live in
// the B operand because rows are the reduction index here).  d(ll)/d(theta) is re-derived per element from the kept
// g1 and p (-> mu), the count and theta by eight otherwise idle warps of the GEMM kernel while the stage they read is
// being multiplied: 3 SFU ops per element, hidden under the kernel's HBM time.
// Replaces, as tc_dw_kernel does, the autograd backward of the decoder heads + likelihood w.r.t. the head weights
// (reference src/scvi/nn/_base_components.py:386-441, src/scvi/distributions/_negative_binomial.py:36-135).
//
// Roofline: the GEMM streams the kept operands once -- 6 parts x B x G x 2 B (ZINB; 198 MB at B = 8192, G = 2000)
// -- so it is HBM-bound (~35 us at 6 TB/s); the tensor work (28 GFLOP in 3-MMA split form) hides under it.
#include "fastmath.cuh"
#include "tc_layout.cuh"

namespace scvi {

enum { XK_F32 = 0, XK_F32_INT = 1, XK_U8 = 2 };
__host__ __device__ inline int64_t dw2_nT4(int64_t G) { return ((G + TC_M - 1) / TC_M) * 4; }
__host__ __device__ inline size_t dw2_part_bytes(int64_t B, int64_t G) {
  return (size_t)((B + TC_M - 1) / TC_M) * 8 * dw2_nT4(G) * 4 * 256;
}

// ---- B operands: per 16-row piece {B1, B2, B3} x {hi, lo} x [18 chunks of 8 columns][16 rows][16 B] (columns 0..127 =
// h, column 128 = 1, the rest 0; each scaled by the piece rows' factors) + a row table [16 rows][8 floats]:
// {1/s1, 2^-P exp(library), u, valid, row id (int64 bits), -, -} for the d(ll)/d(theta) warps ----------------------
constexpr uint32_t DW2_BV = 18 * 256;                 // one variant part
constexpr uint32_t DW2_RT = 16 * 32;                  // row table
constexpr uint32_t DW2_BSTAGE = 6 * DW2_BV + DW2_RT;  // 28,160 B

// scal[0..1] = 2^k12, 2^k3 (so that the largest row factor of each operand maps into (0.5, 1])
__global__ void __launch_bounds__(512)
    tc_pack_b16_kernel(const float* __restrict__ h, int64_t ldh, int64_t B, const float* __restrict__ row_stats,
                       const float* __restrict__ library, const float* __restrict__ g_reconst,
                       const int64_t* __restrict__ row_idx, const float* __restrict__ fmax_part, int n_fmax,
                       uint8_t* __restrict__ bpack, float* __restrict__ scal) {
  __shared__ float s_m[3];
  __shared__ float s_f[TC_M][4];
  const int tid = threadIdx.x;
  if (tid < 96) {  // fixed-order max of the block maxima (a max: any order gives the same value)
    const int q = tid >> 5, lane = tid & 31;
    float m = 0.f;
    for (int i = lane; i < n_fmax; i += 32) m = fmaxf(m, fmax_part[(int64_t)i * 3 + q]);
    m = warp_max(m);
    if (lane == 0) s_m[q] = m;
  }
  __syncthreads();
  auto k_of = [](float m) { return -max(-100, min(100, (int)ceilf(log2f(fmaxf(m, 1e-30f))))); };
  const int k12 = k_of(fmaxf(s_m[0], s_m[1]));
  const int k3 = k_of(s_m[2] * tc::pow2i(-TC_DP_SCALE_LOG2));
  if (blockIdx.x == 0 && tid == 0) {
    scal[0] = tc::pow2i(k12);
    scal[1] = tc::pow2i(k3);
  }
  const int64_t r0 = (int64_t)blockIdx.x * TC_M;
  if (tid < TC_M) {
    const int64_t b = r0 + tid;
    float f1 = 0.f, f2 = 0.f, f3 = 0.f;
    float4 t0 = make_float4(0.f, 0.f, 0.f, 0.f), t1 = t0;
    if (b < B) {
      const float u = g_reconst ? g_reconst[b] : 1.f;
      const float S = row_stats[b * 4 + 3];
      const float lib = library[b];
      const int e = max(-100, min(100, (int)ceilf(lib * fm::LOG2E)));  // K2's operand scale s1 = 2^(14 - e)
      f1 = u * tc::pow2i(e - 14) * tc::pow2i(k12);
      f2 = -u * S * tc::pow2i(-TC_P_SCALE_LOG2) * tc::pow2i(k12);
      f3 = u * tc::pow2i(-TC_DP_SCALE_LOG2) * tc::pow2i(k3);
      const int64_t rid = row_idx ? row_idx[b] : b;
      t0 = make_float4(tc::pow2i(e - 14), expf(lib) * tc::pow2i(-TC_P_SCALE_LOG2), u, 1.f);
      t1 = make_float4(__uint_as_float((uint32_t)(rid & 0xffffffffu)), __uint_as_float((uint32_t)((uint64_t)rid >> 32)), 0.f, 0.f);
    }
    s_f[tid][0] = f1; s_f[tid][1] = f2; s_f[tid][2] = f3; s_f[tid][3] = 0.f;
    float* rt = reinterpret_cast<float*>(bpack + ((int64_t)blockIdx.x * 8 + (tid >> 4)) * DW2_BSTAGE + 6 * DW2_BV) + (tid & 15) * 8;
    *reinterpret_cast<float4*>(rt) = t0;
    *reinterpret_cast<float4*>(rt + 4) = t1;
  }
  __syncthreads();
  for (int it = tid; it < TC_M * 18; it += 512) {
    const int r = it % TC_M, c = it / TC_M;  // chunk of the B variants
    const int64_t b = r0 + r;
    uint8_t* piece = bpack + ((int64_t)blockIdx.x * 8 + (r >> 4)) * DW2_BSTAGE + (r & 15) * 16;
    float hs[8];
    if (c < 16 && b < B) {
      const float4 a = *reinterpret_cast<const float4*>(h + b * ldh + 8 * c);
      const float4 d = *reinterpret_cast<const float4*>(h + b * ldh + 8 * c + 4);
      hs[0] = a.x; hs[1] = a.y; hs[2] = a.z; hs[3] = a.w; hs[4] = d.x; hs[5] = d.y; hs[6] = d.z; hs[7] = d.w;
#pragma unroll
      for (int q = 0; q < 8; ++q) hs[q] = fminf(fmaxf(hs[q] * TC_SCALE_H, -60000.f), 60000.f);
    } else {
#pragma unroll
      for (int q = 0; q < 8; ++q) hs[q] = 0.f;
      if (c == 16 && b < B) hs[0] = TC_SCALE_H;  // the ones column, in the unit of the h columns
    }
#pragma unroll
    for (int v = 0; v < 3; ++v) {
      float w[8];
#pragma unroll
      for (int q = 0; q < 8; ++q) w[q] = hs[q] * s_f[r][v];
      uint4 hi, lo;
      tc::split8(w, hi, lo);
      *reinterpret_cast<uint4*>(piece + (2 * v) * DW2_BV + c * 256) = hi;
      *reinterpret_cast<uint4*>(piece + (2 * v + 1) * DW2_BV + c * 256) = lo;
    }
  }
}

// ---- the GEMM --------------------------------------------------------------------------------------------------
constexpr int DW2_NS = 3;        // stages
constexpr int DW2_DTW = 16;      // d(ll)/d(theta) warps: warps 4..19 (thread = gene, four quarters of the stage's 16 rows)
constexpr int DW2_RPW = 16 / (DW2_DTW / 4);  // rows of a stage per warp group
constexpr int DW2_THREADS = 32 * (4 + DW2_DTW);
constexpr int DW2_T = 32;        // digamma-difference table length per gene
template <bool ZINB>
struct Dw2Smem {
  static constexpr int NP = ZINB ? 6 : 4;
  static constexpr uint32_t A_BYTES = NP * 4096;
  static constexpr uint32_t STAGE = A_BYTES + DW2_BSTAGE;
  static constexpr uint32_t BARS = DW2_NS * STAGE;
  static constexpr uint32_t TMEM_PTR = BARS + 16 * 8;
  static constexpr uint32_t DGT = TMEM_PTR + 16;                  // [DW2_T][128 genes] psi(k + theta) - psi(theta)
  static constexpr uint32_t RED = DGT + DW2_T * TC_M * 4;          // [3][128] partial sums of the other row quarters
  static constexpr uint32_t TOTAL = RED + 3 * TC_M * 4;
};
static_assert(Dw2Smem<true>::TOTAL + 128 <= 232448, "dW GEMM shared memory");

// CTA = (128-gene block gb, split of the 16-row pieces).  TMEM: D12 [128 genes x 144] at 0, D3 at 144.
template <bool ZINB, int XK>
__global__ void __launch_bounds__(DW2_THREADS, 1)
    tc_dw_gemm_kernel(int64_t B, int64_t G, const uint8_t* __restrict__ aops, const uint8_t* __restrict__ bpack,
                      const float* __restrict__ scal, const void* __restrict__ xv, int64_t ldx,
                      const float* __restrict__ px_r, float eps, int pieces_per_split,
                      float* __restrict__ slabW /*[splits][NH][G][H]*/, float* __restrict__ slabv /*[splits][3][G]*/) {
  using L = Dw2Smem<ZINB>;
  constexpr int NH = ZINB ? 2 : 1;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((128u - (tc::smem_u32(smem_raw) & 127u)) & 127u);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::BARS);
  uint64_t *full = bars, *empty = bars + DW2_NS, *accfull = bars + 2 * DW2_NS;
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(smem + L::TMEM_PTR);
  float* s_dgt = reinterpret_cast<float*>(smem + L::DGT);
  float* s_red = reinterpret_cast<float*>(smem + L::RED);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int64_t gb = blockIdx.x;
  const int64_t nP = ((B + TC_M - 1) / TC_M) * 8;  // 16-row pieces
  const int64_t p0 = (int64_t)blockIdx.y * pieces_per_split;
  const int n = (int)max((int64_t)0, min(nP, p0 + pieces_per_split) - p0);
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < DW2_NS; ++i) {
      tc::mbar_init(&full[i], 1);
      tc::mbar_init(&empty[i], 1 + DW2_DTW);  // the MMA commit + the d(ll)/d(theta) warps that read the stage
    }
    tc::mbar_init(accfull, 1);
    tc::fence_mbar_init();
  }
  if (warp == 2) tc::tmem_alloc(s_tmem, 512);
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem_base = *s_tmem;
  float dth_sum = 0.f;  // (d(ll)/d(theta) warps) sum_b u_b d(ll)/d(theta) of this thread's gene over its rows
  if (warp == 0) {
    // ===== producer: the piece's operand blocks (one 4 KB bulk copy per part) + its B block
    if (lane == 0) {
      const size_t ps = dw2_part_bytes(B, G);
      const int64_t nT4 = dw2_nT4(G);
      for (int i = 0; i < n; ++i) {
        const int s = i % DW2_NS;
        tc::mbar_wait(&empty[s], ((i / DW2_NS) & 1) ^ 1);
        tc::mbar_arrive_expect_tx(&full[s], L::STAGE);
        uint8_t* dst = smem + s * L::STAGE;
        const size_t off = (((size_t)(p0 + i) * nT4 + gb * 4) * 4) * 256;
#pragma unroll
        for (int p = 0; p < L::NP; ++p) tc::bulk_g2s(dst + p * 4096, aops + p * ps + off, 4096, &full[s]);
        tc::bulk_g2s(dst + L::A_BYTES, bpack + (size_t)(p0 + i) * DW2_BSTAGE, DW2_BSTAGE, &full[s]);
      }
    }
  } else if (warp == 1) {
    if (n > 0) {  // warp-uniform issue loop, one elected lane issues (tc.cuh)
      constexpr uint32_t idW = tc::idesc_f16_f32(TC_M, 144, 1, 1);
      const uint32_t leader = tc::elect_one();
      const uint32_t s0 = tc::smem_u32(smem);
      // MN-major blocks [chunk of 8][16 rows][16 B]: LBO (next 8 reduction rows) = 128, SBO (next chunk) = 256
      auto opA = [&](uint32_t base, int p) { return tc::make_oper(base + p * 4096, base + (p + 1) * 4096, 128, 256, 0); };
      auto opB = [&](uint32_t base, int v) {
        return tc::make_oper(base + L::A_BYTES + (2 * v) * DW2_BV, base + L::A_BYTES + (2 * v + 1) * DW2_BV, 128, 256, 0);
      };
      for (int i = 0; i < n; ++i) {
        const int s = i % DW2_NS;
        tc::mbar_wait(&full[s], (i / DW2_NS) & 1);
        tc::fence_after_sync();
        if (leader) {
          const uint32_t base = s0 + s * L::STAGE;
          tc::mma3<1>(tmem_base, opA(base, 0), opB(base, 0), idW, i > 0);
          tc::mma3<1>(tmem_base, opA(base, 2), opB(base, 1), idW, true);
          if (ZINB) tc::mma3<1>(tmem_base + 144, opA(base, 4), opB(base, 2), idW, i > 0);
          tc::mma_commit(&empty[s]);
        }
        __syncwarp();
      }
      if (leader) tc::mma_commit(accfull);
      __syncwarp();
    }
  } else if (warp >= 4) {
    // ===== d(ll)/d(theta): thread = gene gl of the block, warp group q = (warp - 4) / 4 takes rows 4q..4q+3 of a stage
    const int gl = 32 * (warp & 3) + lane, half = (warp - 4) >> 2;
    const int64_t g = gb * TC_M + gl;
    const bool gvalid = g < G;
    const int64_t gx = min(g, G - 1);  // a valid column to read for the threads past G
    const float theta = expf(px_r[gx]);
    const float rth = 1.f / (theta + eps);
    if (half == 0) {  // psi(k + theta) - psi(theta) = sum_{i < k} 1 / (theta + i)
      float d = 0.f;
      for (int k = 0; k < DW2_T; ++k) {
        s_dgt[k * TC_M + gl] = d;
        d += __frcp_rn(theta + (float)k);
      }
    }
    asm volatile("bar.sync 1, %0;" ::"n"(32 * DW2_DTW) : "memory");
    const uint32_t eoff = (uint32_t)(gl >> 3) * 256 + (gl & 7) * 2 + half * (DW2_RPW * 16);  // (chunk, first row of the half, element)
    // counts of this thread's 8 (row, gene) elements, software-pipelined two deep: the row ids of piece i + 2 and the
    // counts of piece i + 1 are loaded (from the B block's row table in global memory / the count matrix) while
    // piece i is processed, so neither global latency sits between `full` and the math
    auto rid_ptr = [&](int i, int r) {
      return reinterpret_cast<const int64_t*>(bpack + (size_t)(p0 + i) * DW2_BSTAGE + 6 * DW2_BV + (half * DW2_RPW + r) * 32 + 16);
    };
    auto load_x = [&](int64_t rid) -> int {
      if (XK == XK_U8) return (int)reinterpret_cast<const uint8_t*>(xv)[rid * ldx + gx];
      return (int)reinterpret_cast<const float*>(xv)[rid * ldx + gx];
    };
    int64_t rid_n[DW2_RPW];
    int xa[DW2_RPW], xb[DW2_RPW];
#pragma unroll
    for (int r = 0; r < DW2_RPW; ++r) {
      xa[r] = n > 0 ? load_x(*rid_ptr(0, r)) : 0;
      rid_n[r] = n > 1 ? *rid_ptr(1, r) : 0;
      xb[r] = 0;
    }
    // one stage: this thread's DW2_RPW elements with the counts `xs`
    auto stage_math = [&](int i, const int (&xs)[DW2_RPW]) {
      const int s = i % DW2_NS;
      tc::mbar_wait(&full[s], (i / DW2_NS) & 1);
      const uint8_t* st = smem + s * L::STAGE;
      const float* rt = reinterpret_cast<const float*>(st + L::A_BYTES + 6 * DW2_BV) + half * DW2_RPW * 8;
#pragma unroll
      for (int r = 0; r < DW2_RPW; ++r) {
        const float4 t0 = *reinterpret_cast<const float4*>(rt + r * 8);  // 1/s1, 2^-P exp(library), u, valid
        const uint8_t* e = st + eoff + r * 16;
        const float a1 = __half2float(*reinterpret_cast<const __half*>(e)) + __half2float(*reinterpret_cast<const __half*>(e + 4096));
        const float a2 = __half2float(*reinterpret_cast<const __half*>(e + 2 * 4096)) +
                         __half2float(*reinterpret_cast<const __half*>(e + 3 * 4096));
        const float g1 = a1 * t0.x, mu = a2 * t0.y;
        const int xi = xs[r];
        const float x = (float)xi;
        const float rT = fm::rcp(theta + mu + eps);
        const float d_log = -fm::LN2 * fm::lg2(fmaf(mu, rth, 1.f));  // log(theta+eps) - log(theta+mu+eps)
        const float da = fmaf(theta, rth - rT, d_log);
        float dgd;
        if (xi < DW2_T) {
          dgd = s_dgt[xi * TC_M + gl];
        } else {
          dgd = s_dgt[(DW2_T - 1) * TC_M + gl];
          for (int k = DW2_T - 1; k < xi; ++k) dgd += __frcp_rn(theta + (float)k);
        }
        float dth = fmaf(-x, rT, da) + dgd;
        if (ZINB && xi == 0) {  // zero-inflated zero: weight sigmoid(a) = g1 / (theta mu / (theta + mu + eps))
          const float base = theta * mu * rT;
          const float sa = base > 1e-30f ? fminf(fmaxf(g1 * fm::rcp(base), 0.f), 1.f) : 0.f;
          dth = sa * da;
        }
        dth_sum = fmaf(t0.z * t0.w, dth, dth_sum);
      }
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(&empty[s]);
    };
    // two stages per trip so the count registers alternate without moves (a move would wait for the loads in flight)
    for (int i = 0; i < n; i += 2) {
      if (i + 1 < n) {
#pragma unroll
        for (int r = 0; r < DW2_RPW; ++r) xb[r] = load_x(rid_n[r]);
      }
      if (i + 2 < n) {
#pragma unroll
        for (int r = 0; r < DW2_RPW; ++r) rid_n[r] = *rid_ptr(i + 2, r);
      }
      stage_math(i, xa);
      if (i + 1 < n) {
        if (i + 2 < n) {
#pragma unroll
          for (int r = 0; r < DW2_RPW; ++r) xa[r] = load_x(rid_n[r]);
        }
        if (i + 3 < n) {
#pragma unroll
          for (int r = 0; r < DW2_RPW; ++r) rid_n[r] = *rid_ptr(i + 3, r);
        }
        stage_math(i + 1, xb);
      }
    }
    if (half > 0) s_red[(half - 1) * TC_M + gl] = dth_sum;
    asm volatile("bar.sync 1, %0;" ::"n"(32 * DW2_DTW) : "memory");
    if (half == 0) {
      dth_sum += (s_red[gl] + s_red[TC_M + gl]) + s_red[2 * TC_M + gl];
      // ===== epilogue (warps 4..7): thread = gene; TMEM lane = gene
      if (n > 0) {
        tc::mbar_wait(accfull, 0);
        tc::fence_after_sync();
      }
      const uint32_t lane_sel = (uint32_t)(32 * (warp & 3)) << 16;
      const float kW = 1.f / (scal[0] * TC_SCALE_H), kP = 1.f / (scal[1] * TC_SCALE_H);
#pragma unroll 1
      for (int hd = 0; hd < NH; ++hd) {
        float* o = slabW + (((int64_t)blockIdx.y * NH + hd) * G + g) * TC_H;
        const float kk = hd == 0 ? kW : kP;
#pragma unroll 1
        for (int c = 0; c < 17; ++c) {  // 16 chunks of weights + the ones-column chunk
          float v[8];
          if (n > 0) {
            tc::tmem_ld8(tmem_base + lane_sel + hd * 144 + 8 * c, v);
            tc::tmem_ld_wait();
          } else {
#pragma unroll
            for (int i = 0; i < 8; ++i) v[i] = 0.f;
          }
          if (!gvalid) continue;
          if (c < 16) {
            *reinterpret_cast<float4*>(o + 8 * c) = make_float4(v[0] * kk, v[1] * kk, v[2] * kk, v[3] * kk);
            *reinterpret_cast<float4*>(o + 8 * c + 4) = make_float4(v[4] * kk, v[5] * kk, v[6] * kk, v[7] * kk);
          } else {
            slabv[((int64_t)blockIdx.y * 3 + hd) * G + g] = v[0] * kk;  // dbs / dbp
          }
        }
      }
      if (gvalid) {
        if (!ZINB) slabv[((int64_t)blockIdx.y * 3 + 1) * G + g] = 0.f;
        slabv[((int64_t)blockIdx.y * 3 + 2) * G + g] = dth_sum;  // sum_b u_b d(ll)/d(theta)
      }
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 2) tc::tmem_dealloc(tmem_base, 512);
}

// ---- fixed-order sums of the split slabs -> gradients -----------------------------------------------------------
static __global__ void tc_dw2_reduce_kernel(const float* __restrict__ slabW, const float* __restrict__ slabv,
                                            int64_t splits, int nheads, int64_t GH, int64_t G, const float* __restrict__ px_r,
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
    if (dpx_r) dpx_r[g] = -expf(px_r[g]) * c;
  }
}

// ---- host -------------------------------------------------------------------------------------------------------
struct Dw2Plan {
  int64_t nGB, nP;
  int splits, pieces_per_split;
  size_t off_bpack, off_scal, off_slabW, off_slabv, total;
};
static Dw2Plan dw2_plan(int64_t B, int64_t G, int nheads) {
  Dw2Plan p;
  p.nGB = cdiv(G, TC_M);
  p.nP = cdiv(B, TC_M) * 8;
  int64_t s = std::min<int64_t>(std::max<int64_t>(1, 148 / p.nGB), p.nP);
  p.pieces_per_split = (int)cdiv(p.nP, s);
  p.splits = (int)cdiv(p.nP, p.pieces_per_split);
  auto al = [](size_t v) { return (v + 255) / 256 * 256; };
  p.off_bpack = 0;
  p.off_scal = al((size_t)p.nP * DW2_BSTAGE);
  p.off_slabW = al(p.off_scal + 64);
  p.off_slabv = al(p.off_slabW + (size_t)p.splits * nheads * G * TC_H * sizeof(float));
  p.total = al(p.off_slabv + (size_t)p.splits * 3 * G * sizeof(float));
  return p;
}
int64_t tc_decoder_bwd2_ws_bytes(int64_t B, int64_t G, int likelihood) {
  return (int64_t)dw2_plan(B, G, likelihood == SCVI_LIK_ZINB ? 2 : 1).total;
}

int tc_decoder_nll_bwd_dw2(const float* h, int64_t ldh, int64_t B, int64_t G, const float* px_r, const void* x,
                           int64_t ldx, const int64_t* row_idx, const float* library, const float* row_stats,
                           const float* g_reconst, int likelihood, float* dWs, float* dbs, float* dWp, float* dbp,
                           float* dpx_r, void* ws, const uint8_t* aops, const float* fmax_part, int xkind,
                           cudaStream_t st) {
  const bool zinb = likelihood == SCVI_LIK_ZINB;
  const int nheads = zinb ? 2 : 1;
  SCVI_REQUIRE(ws != nullptr && ((uintptr_t)ws % 256 == 0), "scvi_decoder_nll_bwd(SCVI_PREC_AOPS): ws must be 256-byte aligned");
  SCVI_REQUIRE(library != nullptr, "scvi_decoder_nll_bwd(SCVI_PREC_AOPS): needs `library`");
  SCVI_REQUIRE(ldh % 4 == 0 && ((uintptr_t)h % 16 == 0), "scvi_decoder_nll_bwd(SCVI_PREC_AOPS): h must be 16-byte aligned");
  const Dw2Plan p = dw2_plan(B, G, nheads);
  uint8_t* w8 = (uint8_t*)ws;
  uint8_t* bpack = w8 + p.off_bpack;
  float* scal = (float*)(w8 + p.off_scal);
  float* slabW = (float*)(w8 + p.off_slabW);
  float* slabv = (float*)(w8 + p.off_slabv);
  tc_pack_b16_kernel<<<(unsigned)cdiv(B, TC_M), 512, 0, st>>>(h, ldh, B, row_stats, library, g_reconst, row_idx, fmax_part,
                                                              (int)cdiv(B, 8), bpack, scal);
  SCVI_LAUNCH_CHECK("scvi_decoder_nll_bwd(tc pack B)");
  dim3 grid((unsigned)p.nGB, (unsigned)p.splits);
#define SCVI_DW2(Z, K)                                                                                          \
  do {                                                                                                          \
    auto k = tc_dw_gemm_kernel<Z, K>;                                                                           \
    constexpr int smem = (int)Dw2Smem<Z>::TOTAL + 128;                                                          \
    static std::atomic<uint64_t> optin{0};                                                                      \
    if (int rc = smem_optin((const void*)k, smem, optin, "tc_dw_gemm")) return rc;                              \
    k<<<grid, DW2_THREADS, smem, st>>>(B, G, aops, bpack, scal, x, ldx, px_r, 1e-8f, p.pieces_per_split, slabW, \
                                       slabv);                                                                  \
  } while (0)
  if (zinb) {
    if (xkind == XK_U8) SCVI_DW2(true, XK_U8);
    else SCVI_DW2(true, XK_F32_INT);
  } else {
    if (xkind == XK_U8) SCVI_DW2(false, XK_U8);
    else SCVI_DW2(false, XK_F32_INT);
  }
#undef SCVI_DW2
  SCVI_LAUNCH_CHECK("scvi_decoder_nll_bwd(tc dW gemm)");
  const int64_t GH = G * TC_H;
  tc_dw2_reduce_kernel<<<(unsigned)cdiv(std::max<int64_t>(GH / 4, G), 256), 256, 0, st>>>(
      slabW, slabv, p.splits, nheads, GH, G, px_r, dWs, dWp, dbs, dbp, dpx_r);
  SCVI_LAUNCH_CHECK("scvi_decoder_nll_bwd(tc reduce)");
  return SCVI_OK;
}

}  // namespace scvi
