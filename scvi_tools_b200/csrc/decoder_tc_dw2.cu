// Weight-gradient half of the decoder backward as a PURE tensor-core GEMM (SCVI_PREC_AOPS).
//
// The recompute kernel (tc_dw_kernel, decoder_tc_train.cu) re-derives every per-element derivative the fused
// likelihood kernel K2 already had in registers: ~100 us of SFU-bound work per step.  With SCVI_PREC_AOPS K2 keeps its
// staged operands (fp16 hi/lo) in the forward workspace, laid out so that the block a GEMM CTA needs is contiguous,
// and the gradients become
//     dWs = (g1 s1)^T (u/s1 . h)  +  (p 2^P)^T (-u S 2^-P . h)          dbs = the same products with a ones column
//     dWp = (-dpi 2^D)^T (u 2^-D . h)                                    dbp likewise
//     dpx_r = -theta ( (dth')^T u  +  sum_b u_b [psi(x_bg + theta_g) - psi(theta_g)] )
// (u = loss weight of the row, S = the row's sum of g1, s1 / 2^P / 2^D = K2's operand scales; the row factors live in
// the B operand because rows are the reduction index here; dth' = d(ll)/d(theta) without its digamma term, which
// depends on the count and theta only and is summed by tc_dpsi_colsum_kernel beside K2).
// Replaces, as tc_dw_kernel does, the autograd backward of the decoder heads + likelihood w.r.t. the head weights
// (reference src/scvi/nn/_base_components.py:386-441, src/scvi/distributions/_negative_binomial.py:36-135).
//
// Roofline: the GEMM streams the kept operands once -- 8 parts x B x G x 2 B (ZINB; 264 MB at B = 8192, G = 2000)
// -- so it is HBM-bound (~45 us at 6 TB/s); the tensor work (28 GFLOP in 3-MMA split form) hides under it.
#include "fastmath.cuh"
#include "tc_layout.cuh"

namespace scvi {

enum { XK_F32 = 0, XK_F32_INT = 1, XK_U8 = 2 };
__host__ __device__ inline int64_t dw2_nT4(int64_t G) { return ((G + TC_M - 1) / TC_M) * 4; }
__host__ __device__ inline size_t dw2_part_bytes(int64_t B, int64_t G) {
  return (size_t)((B + TC_M - 1) / TC_M) * 8 * dw2_nT4(G) * 4 * 256;
}

// ---- B operands: per 16-row piece {B1, B2, B3} x {hi, lo} x [18 chunks of 8 columns][16 rows][16 B] (columns 0..127 =
// h, column 128 = 1, the rest 0; each scaled by the piece rows' factors) + Bf {hi, lo} x [2 chunks][16 rows][16 B]
// (column 0 = u 2^k4) ------------------------------------------------------------------------------------------------
constexpr uint32_t DW2_BV = 18 * 256;                 // one variant part
constexpr uint32_t DW2_BF = 2 * 256;
constexpr uint32_t DW2_BSTAGE = 6 * DW2_BV + 2 * DW2_BF;  // 28,672 B

// scal[0..2] = 2^k12, 2^k3, 2^k4 (so that the largest row factor of each operand maps into (0.5, 1])
__global__ void __launch_bounds__(512)
    tc_pack_b16_kernel(const float* __restrict__ h, int64_t ldh, int64_t B, const float* __restrict__ row_stats,
                       const float* __restrict__ library, const float* __restrict__ g_reconst,
                       const float* __restrict__ fmax_part, int n_fmax, uint8_t* __restrict__ bpack,
                       float* __restrict__ scal) {
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
  const int k4 = k_of(s_m[2]);
  if (blockIdx.x == 0 && tid == 0) {
    scal[0] = tc::pow2i(k12);
    scal[1] = tc::pow2i(k3);
    scal[2] = tc::pow2i(k4);
  }
  const int64_t r0 = (int64_t)blockIdx.x * TC_M;
  if (tid < TC_M) {
    const int64_t b = r0 + tid;
    float f1 = 0.f, f2 = 0.f, f3 = 0.f, f4 = 0.f;
    if (b < B) {
      const float u = g_reconst ? g_reconst[b] : 1.f;
      const float S = row_stats[b * 4 + 3];
      const int e = max(-100, min(100, (int)ceilf(library[b] * fm::LOG2E)));  // K2's operand scale s1 = 2^(14 - e)
      f1 = u * tc::pow2i(e - 14) * tc::pow2i(k12);
      f2 = -u * S * tc::pow2i(-TC_P_SCALE_LOG2) * tc::pow2i(k12);
      f3 = u * tc::pow2i(-TC_DP_SCALE_LOG2) * tc::pow2i(k3);
      f4 = u * tc::pow2i(k4);
    }
    s_f[tid][0] = f1; s_f[tid][1] = f2; s_f[tid][2] = f3; s_f[tid][3] = f4;
  }
  __syncthreads();
  for (int it = tid; it < TC_M * 19; it += 512) {
    const int r = it % TC_M, c = it / TC_M;  // c < 18: chunk of the B variants; c == 18: the Bf chunks
    const int64_t b = r0 + r;
    uint8_t* piece = bpack + ((int64_t)blockIdx.x * 8 + (r >> 4)) * DW2_BSTAGE + (r & 15) * 16;
    if (c == 18) {
      float v[8] = {s_f[r][3], 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      uint4 hi, lo;
      tc::split8(v, hi, lo);
      uint8_t* o = piece + 6 * DW2_BV;
      *reinterpret_cast<uint4*>(o) = hi;
      *reinterpret_cast<uint4*>(o + 256) = make_uint4(0u, 0u, 0u, 0u);
      *reinterpret_cast<uint4*>(o + DW2_BF) = lo;
      *reinterpret_cast<uint4*>(o + DW2_BF + 256) = make_uint4(0u, 0u, 0u, 0u);
      continue;
    }
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
constexpr int DW2_NS = 3;       // stages
constexpr int DW2_THREADS = 256;
template <bool ZINB>
struct Dw2Smem {
  static constexpr int NP = ZINB ? 8 : 6;
  static constexpr uint32_t A_BYTES = NP * 4096;
  static constexpr uint32_t STAGE = A_BYTES + DW2_BSTAGE;
  static constexpr uint32_t BARS = DW2_NS * STAGE;
  static constexpr uint32_t TMEM_PTR = BARS + 16 * 8;
  static constexpr uint32_t TOTAL = TMEM_PTR + 16;
};
static_assert(Dw2Smem<true>::TOTAL + 128 <= 232448, "dW GEMM shared memory");

// CTA = (128-gene block gb, split of the 16-row pieces).  TMEM: D12 [128 genes x 144] at 0, D3 at 144, D4 [x 16] at 288.
template <bool ZINB>
__global__ void __launch_bounds__(DW2_THREADS, 1)
    tc_dw_gemm_kernel(int64_t B, int64_t G, const uint8_t* __restrict__ aops, const uint8_t* __restrict__ bpack,
                      const float* __restrict__ scal, int pieces_per_split, float* __restrict__ slabW /*[splits][NH][G][H]*/,
                      float* __restrict__ slabv /*[splits][3][G]*/) {
  using L = Dw2Smem<ZINB>;
  constexpr int NH = ZINB ? 2 : 1;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((128u - (tc::smem_u32(smem_raw) & 127u)) & 127u);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::BARS);
  uint64_t *full = bars, *empty = bars + DW2_NS, *accfull = bars + 2 * DW2_NS;
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(smem + L::TMEM_PTR);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int64_t gb = blockIdx.x;
  const int64_t nP = ((B + TC_M - 1) / TC_M) * 8;  // 16-row pieces
  const int64_t p0 = (int64_t)blockIdx.y * pieces_per_split;
  const int n = (int)max((int64_t)0, min(nP, p0 + pieces_per_split) - p0);
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < DW2_NS; ++i) {
      tc::mbar_init(&full[i], 1);
      tc::mbar_init(&empty[i], 1);
    }
    tc::mbar_init(accfull, 1);
    tc::fence_mbar_init();
  }
  if (warp == 2) tc::tmem_alloc(s_tmem, 512);
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem_base = *s_tmem;
  if (warp == 0) {
    // ===== producer: the piece's operand blocks (one 4 KB bulk copy per part) + its 28 KB B block
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
      constexpr uint32_t idF = tc::idesc_f16_f32(TC_M, 16, 1, 1);
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
          if (ZINB) tc::mma3<1>(tmem_base + 144, opA(base, 6), opB(base, 2), idW, i > 0);
          const tc::Oper bf = tc::make_oper(base + L::A_BYTES + 6 * DW2_BV, base + L::A_BYTES + 6 * DW2_BV + DW2_BF, 128, 256, 0);
          tc::mma3<1>(tmem_base + 288, opA(base, 4), bf, idF, i > 0);
          tc::mma_commit(&empty[s]);
        }
        __syncwarp();
      }
      if (leader) tc::mma_commit(accfull);
      __syncwarp();
    }
  } else if (warp >= 4) {
    // ===== epilogue: thread = gene 32 * (warp % 4) + lane of the block
    const int gl = 32 * (warp & 3) + lane;
    const int64_t g = gb * TC_M + gl;
    const bool gvalid = g < G;
    if (n > 0) {
      tc::mbar_wait(accfull, 0);
      tc::fence_after_sync();
    }
    const uint32_t lane_sel = (uint32_t)(32 * (warp & 3)) << 16;
    const float kW = 1.f / (scal[0] * TC_SCALE_H), kP = 1.f / (scal[1] * TC_SCALE_H), kT = 1.f / scal[2];
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
    float v[8];
    if (n > 0) {
      tc::tmem_ld8(tmem_base + lane_sel + 288, v);
      tc::tmem_ld_wait();
    } else {
      v[0] = 0.f;
    }
    if (gvalid) {
      if (!ZINB) slabv[((int64_t)blockIdx.y * 3 + 1) * G + g] = 0.f;
      slabv[((int64_t)blockIdx.y * 3 + 2) * G + g] = v[0] * kT;  // sum_b u dth'
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 2) tc::tmem_dealloc(tmem_base, 512);
}

// ---- sum_b u_b [psi(x_bg + theta_g) - psi(theta_g)] per gene: thread = gene, rows of one split in order.
// psi(x + theta) - psi(theta) = sum_{i < x} 1 / (theta + i) for an integer count x: a 32-entry table per gene in shared
// memory ([count][gene]: conflict-free), the rare larger counts by the series / asymptotic digamma.  Depends on the
// counts, theta and the loss weights only: the training plan runs it on its side stream beside the forward.
constexpr int DPSI_T = 32;
template <int XK>
__global__ void __launch_bounds__(128)
    tc_dpsi_colsum_kernel(const void* __restrict__ xv, int64_t ldx, const int64_t* __restrict__ row_idx, int64_t B,
                          int64_t G, const float* __restrict__ px_r, const float* __restrict__ g_reconst,
                          int rows_per_split, float* __restrict__ slab /*[splits][G]*/) {
  __shared__ float s_T[DPSI_T * 128];
  __shared__ int64_t s_row[64];
  __shared__ float s_u[64];
  const int64_t g = (int64_t)blockIdx.x * 128 + threadIdx.x;
  const int64_t b0 = (int64_t)blockIdx.y * rows_per_split, b1 = min(B, b0 + rows_per_split);
  const float theta = g < G ? expf(px_r[g]) : 1.f;
  {
    float d = 0.f;
    for (int k = 0; k < DPSI_T; ++k) {
      s_T[k * 128 + threadIdx.x] = d;
      d += __frcp_rn(theta + (float)k);
    }
  }
  const int64_t gc = min(g, G - 1);  // threads past G read a valid column and drop the result
  float acc = 0.f;
  for (int64_t bb = b0; bb < b1; bb += 64) {
    __syncthreads();
    if (threadIdx.x < 64 && bb + threadIdx.x < b1) {
      const int64_t b = bb + threadIdx.x;
      s_row[threadIdx.x] = row_idx ? row_idx[b] : b;
      s_u[threadIdx.x] = g_reconst ? g_reconst[b] : 1.f;
    }
    __syncthreads();
    const int nrow = (int)min((int64_t)64, b1 - bb);
    for (int rb = 0; rb < nrow; rb += 16) {
      int xs[16];  // the 16 loads are issued together (no branch between them), then consumed
#pragma unroll
      for (int k = 0; k < 16; ++k) {
        const int r = min(rb + k, nrow - 1);
        if (XK == XK_U8) xs[k] = (int)reinterpret_cast<const uint8_t*>(xv)[s_row[r] * ldx + gc];
        else xs[k] = (int)reinterpret_cast<const float*>(xv)[s_row[r] * ldx + gc];
      }
#pragma unroll
      for (int k = 0; k < 16; ++k) {
        if (rb + k >= nrow) break;
        const int xi = xs[k];
        float d;
        if (xi < DPSI_T) {
          d = s_T[xi * 128 + threadIdx.x];
        } else {
          d = s_T[(DPSI_T - 1) * 128 + threadIdx.x];
          if (xi < 256) {
            for (int i = DPSI_T - 1; i < xi; ++i) d += __frcp_rn(theta + (float)i);
          } else {
            d = fm::digamma_pos_fast((float)xi + theta) - fm::digamma_pos_fast(theta);
          }
        }
        acc = fmaf(s_u[rb + k], d, acc);
      }
    }
  }
  if (g < G) slab[(int64_t)blockIdx.y * G + g] = acc;
}

// ---- fixed-order sums of the split slabs -> gradients -----------------------------------------------------------
static __global__ void tc_dw2_reduce_kernel(const float* __restrict__ slabW, const float* __restrict__ slabv,
                                            const float* __restrict__ slabpsi, int64_t splits, int64_t splits_psi,
                                            int nheads, int64_t GH, int64_t G, const float* __restrict__ px_r,
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
    float a = 0.f, b = 0.f, c = 0.f, d = 0.f;
#pragma unroll 4
    for (int64_t t = 0; t < splits; ++t) {
      a += slabv[(t * 3 + 0) * G + g];
      b += slabv[(t * 3 + 1) * G + g];
      c += slabv[(t * 3 + 2) * G + g];
    }
    for (int64_t t = 0; t < splits_psi; ++t) d += slabpsi[t * G + g];
    if (dbs) dbs[g] = a;
    if (nheads == 2 && dbp) dbp[g] = b;
    if (dpx_r) dpx_r[g] = -expf(px_r[g]) * (c + d);
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

// The count-only digamma sums (SCVI_PREC_AOPS): [dpsi_splits(B)][G] floats at `slab`
int dpsi_splits(int64_t B) { return (int)cdiv(B, cdiv(B, std::max<int64_t>(1, std::min<int64_t>(32, B / 256)))); }
size_t tc_dpsi_bytes(int64_t B, int64_t G) { return (size_t)dpsi_splits(B) * G * sizeof(float); }
int tc_dpsi_sums(const void* x, int64_t ldx, const int64_t* row_idx, int64_t B, int64_t G, const float* px_r,
                 const float* g_reconst, int xkind, float* slab, cudaStream_t st) {
  const int splits = dpsi_splits(B);
  const int rows_per_split = (int)cdiv(B, splits);
  dim3 grid((unsigned)cdiv(G, 128), (unsigned)splits);
  if (xkind == XK_U8)
    tc_dpsi_colsum_kernel<XK_U8><<<grid, 128, 0, st>>>(x, ldx, row_idx, B, G, px_r, g_reconst, rows_per_split, slab);
  else
    tc_dpsi_colsum_kernel<XK_F32_INT><<<grid, 128, 0, st>>>(x, ldx, row_idx, B, G, px_r, g_reconst, rows_per_split, slab);
  SCVI_LAUNCH_CHECK("scvi_decoder_dpsi_sums");
  return SCVI_OK;
}

int tc_decoder_nll_bwd_dw2(const float* h, int64_t ldh, int64_t B, int64_t G, const float* px_r, const void* x,
                           int64_t ldx, const int64_t* row_idx, const float* library, const float* row_stats,
                           const float* g_reconst, int likelihood, float* dWs, float* dbs, float* dWp, float* dbp,
                           float* dpx_r, void* ws, const uint8_t* aops, const float* fmax_part, const float* slabpsi,
                           cudaStream_t st) {
  (void)x; (void)ldx; (void)row_idx;
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
  tc_pack_b16_kernel<<<(unsigned)cdiv(B, TC_M), 512, 0, st>>>(h, ldh, B, row_stats, library, g_reconst, fmax_part,
                                                              (int)cdiv(B, 8), bpack, scal);
  SCVI_LAUNCH_CHECK("scvi_decoder_nll_bwd(tc pack B)");
  dim3 grid((unsigned)p.nGB, (unsigned)p.splits);
  if (zinb) {
    auto k = tc_dw_gemm_kernel<true>;
    constexpr int smem = (int)Dw2Smem<true>::TOTAL + 128;
    static std::atomic<uint64_t> optin{0};
    if (int rc = smem_optin((const void*)k, smem, optin, "tc_dw_gemm")) return rc;
    k<<<grid, DW2_THREADS, smem, st>>>(B, G, aops, bpack, scal, p.pieces_per_split, slabW, slabv);
  } else {
    auto k = tc_dw_gemm_kernel<false>;
    constexpr int smem = (int)Dw2Smem<false>::TOTAL + 128;
    static std::atomic<uint64_t> optin{0};
    if (int rc = smem_optin((const void*)k, smem, optin, "tc_dw_gemm")) return rc;
    k<<<grid, DW2_THREADS, smem, st>>>(B, G, aops, bpack, scal, p.pieces_per_split, slabW, slabv);
  }
  SCVI_LAUNCH_CHECK("scvi_decoder_nll_bwd(tc dW gemm)");
  const int64_t GH = G * TC_H;
  tc_dw2_reduce_kernel<<<(unsigned)cdiv(std::max<int64_t>(GH / 4, G), 256), 256, 0, st>>>(
      slabW, slabv, slabpsi, p.splits, dpsi_splits(B), nheads, GH, G, px_r, dWs, dWp, dbs, dbp, dpx_r);
  SCVI_LAUNCH_CHECK("scvi_decoder_nll_bwd(tc reduce)");
  return SCVI_OK;
}

}  // namespace scvi
