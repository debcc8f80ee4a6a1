// tcgen05 / TMEM variant of the fused decoder heads + softmax + NB/ZINB likelihood (SCVI_PREC_TC).
//
// GEMM: logits[128 rows, 128 genes] = h[128, H=128] * W_tile[128 genes, 128]^T on the 5th-gen tensor
// cores (tcgen05.mma kind::f16, fp32 accumulators in TMEM).  fp32 parity is kept with an
// error-compensated fp16 split: v = hi + lo (2 x 11-bit significands), D = hi*hi + hi*lo + lo*hi
// (3 MMAs per K-step, error ~2^-22).  Operands are pre-scaled by powers of two (h x 4, W x 256) so the
// `lo` parts stay in fp16's normal range; the accumulator is rescaled by 2^-10 in the epilogue.
//
// Pipeline per CTA (128 rows x a contiguous range of 128-gene tiles), warp-specialised:
//   warp 0     : producer   -- 1-D bulk async copies (TMA engine) of pre-packed W tiles, 2-stage ring
//   warp 1     : MMA issuer -- one elected lane issues 24 tcgen05.mma per head tile, commits to mbarriers
//   warp 2     : TMEM allocator
//   warps 4-11 : epilogue   -- tcgen05.ld 16 columns at a time, per-element likelihood, row reductions
// TMEM holds 2 buffers x (1|2 heads) x 128 fp32 columns so the MMAs of tile j+1 overlap the epilogue of j.
#include "decoder_math.cuh"
#include "tc_layout.cuh"

namespace scvi {

constexpr int TC_EPI_WARPS = 16;
constexpr int TC_THREADS = 32 * (4 + TC_EPI_WARPS);

// ---------------------------------------------------------------------------------------
// pack: W fp32 [G, H] -> fp16 hi / lo blocks in the UMMA no-swizzle layout (tc_layout.cuh), twice:
//   packed128: per (128-gene tile, head) {hi block, lo block} of 128 x 128   (LSE / NLL-only / dW kernels)
//   packed32 : per 32-gene tile {Ws hi, Ws lo, (Wp hi, Wp lo), gene constants}  (fused NLL + dh kernel)
// plus the per-gene constant table gconst [nT128 * 128][TC_GC].
// ---------------------------------------------------------------------------------------
__global__ void tc_pack_heads_kernel(const float* __restrict__ Ws, const float* __restrict__ bs,
                                     const float* __restrict__ Wp, const float* __restrict__ bp,
                                     const float* __restrict__ px_r, int64_t G, int nheads, float eps,
                                     __half* __restrict__ packed, uint8_t* __restrict__ packed32,
                                     float* __restrict__ gconst) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t nT = (G + TC_N - 1) / TC_N;
  const int64_t items = nT * nheads * TC_KC * TC_N;
  const uint32_t st32 = tc_w32_stage_bytes(nheads);
  if (e < items) {
    const int r = (int)(e % TC_N);
    const int kc = (int)((e / TC_N) % TC_KC);
    const int head = (int)((e / (TC_N * TC_KC)) % nheads);
    const int64_t t = e / ((int64_t)TC_N * TC_KC * nheads);
    const int64_t g = t * TC_N + r;
    const float* W = head == 0 ? Ws : Wp;
    __half hi[8], lo[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      float v = (g < G) ? W[g * TC_H + 8 * kc + q] * TC_SCALE_W : 0.f;
      v = fminf(fmaxf(v, -60000.f), 60000.f);
      tc::split_f16(v, hi[q], lo[q]);
    }
    if (packed) {
      __half* base = packed + ((t * nheads + head) * 2) * (int64_t)(TC_M * TC_H);
      *reinterpret_cast<uint4*>(base + (int64_t)kc * TC_N * 8 + r * 8) = *reinterpret_cast<uint4*>(hi);
      *reinterpret_cast<uint4*>(base + (int64_t)TC_M * TC_H + (int64_t)kc * TC_N * 8 + r * 8) = *reinterpret_cast<uint4*>(lo);
    }
    if (packed32) {
      const int64_t t32 = g / TC_NT;
      const int r32 = (int)(g % TC_NT);
      uint8_t* base = packed32 + t32 * st32 + (size_t)head * 2 * TC_BLK32;
      *reinterpret_cast<uint4*>(base + (size_t)kc * TC_NT * 16 + r32 * 16) = *reinterpret_cast<uint4*>(hi);
      *reinterpret_cast<uint4*>(base + TC_BLK32 + (size_t)kc * TC_NT * 16 + r32 * 16) = *reinterpret_cast<uint4*>(lo);
    }
  }
  if (e < nT * TC_N) {
    // genes beyond G (tile padding): theta 1, dropout bias 0 and a scale bias of -1e30, so that mu = 0 and every
    // per-element output of the training kernels (ll, g1, dpi, softmax p) is an exact zero without any masking
    const bool ok = e < G;
    const float theta = ok ? expf(px_r[e]) : 1.f;
    float o[TC_GC];
    o[0] = theta;
    o[1] = logf(theta + eps);
    o[2] = lgammaf(theta);
    o[3] = ok ? bs[e] : -1e30f;
    o[4] = (ok && nheads == 2) ? bp[e] : 0.f;
    o[5] = 1.f / (theta + eps);
    o[6] = digamma_pos(theta);
    o[7] = 0.f;
    if (gconst) {
#pragma unroll
      for (int q = 0; q < TC_GC; ++q) gconst[e * TC_GC + q] = o[q];
      gconst[nT * TC_N * TC_GC + e] = o[3];  // contiguous copy of the scale bias for the LSE pass
    }
    if (packed32) {
      float* gc32 = reinterpret_cast<float*>(packed32 + (e / TC_NT) * st32 + (size_t)nheads * 2 * TC_BLK32) + (e % TC_NT) * TC_GC;
#pragma unroll
      for (int q = 0; q < TC_GC; ++q) gc32[q] = o[q];
    }
  }
  // dgd tables for the dW kernel: digamma(x + theta) - digamma(theta) = sum_{k<x} 1/(theta + k), x < DGT, row
  // stride DGT_LD, stored behind the gene constants and the contiguous bias copy
  if (gconst && e < nT * TC_N * fm::DGT_LD) {
    const int64_t g = e / fm::DGT_LD;
    const int xx = (int)(e % fm::DGT_LD);
    const double theta = (g < G) ? (double)expf(px_r[g]) : 1.0;
    double acc = 0.0;
    for (int k = 0; k < xx && k < fm::DGT; ++k) acc += 1.0 / (theta + (double)k);
    gconst[nT * TC_N * (TC_GC + 1) + e] = (float)acc;
  }
  // lgd tables: lgamma(x + theta) - lgamma(theta) - lgamma(x + 1), x < LGT, per gene (double precision)
  if (packed32 && e < nT * TC_N * fm::LGT) {
    const int64_t g = e / fm::LGT;
    const int xx = (int)(e % fm::LGT);
    const double theta = (g < G) ? (double)expf(px_r[g]) : 1.0;
    const float v = xx == 0 ? 0.f : (float)(lgamma((double)xx + theta) - lgamma(theta) - lgamma((double)xx + 1.0));
    float* tab = reinterpret_cast<float*>(packed32 + (g / TC_NT) * st32 + (size_t)nheads * 2 * TC_BLK32 + TC_GC32_BYTES);
    tab[(g % TC_NT) * fm::LGT + xx] = v;
  }
}

// h tile: fp32 [rows, H] -> scaled fp16 hi / lo blocks in shared memory (same layout as the packed W)
__device__ __forceinline__ void tc_stage_h(const float* __restrict__ h, int64_t ldh, int valid, __half* s_hi,
                                           __half* s_lo, int tid, int nthreads) {
  for (int e = tid; e < TC_M * TC_KC; e += nthreads) {
    const int r = e % TC_M, kc = e / TC_M;
    __half hi[8], lo[8];
    if (r < valid) {
      const float4 a = *reinterpret_cast<const float4*>(h + (int64_t)r * ldh + 8 * kc);
      const float4 b = *reinterpret_cast<const float4*>(h + (int64_t)r * ldh + 8 * kc + 4);
      const float v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
      for (int q = 0; q < 8; ++q) tc::split_f16(fminf(fmaxf(v[q] * TC_SCALE_H, -60000.f), 60000.f), hi[q], lo[q]);
    } else {
#pragma unroll
      for (int q = 0; q < 8; ++q) hi[q] = lo[q] = __float2half_rn(0.f);
    }
    *reinterpret_cast<uint4*>(s_hi + kc * TC_M * 8 + r * 8) = *reinterpret_cast<uint4*>(hi);
    *reinterpret_cast<uint4*>(s_lo + kc * TC_M * 8 + r * 8) = *reinterpret_cast<uint4*>(lo);
  }
}

struct TcSmem {
  // dynamic smem carve-up (byte offsets from a 1024-aligned base)
  static constexpr uint32_t H_HI = 0;
  static constexpr uint32_t H_LO = TC_OPER_BYTES;
  static constexpr uint32_t STAGE0 = 2 * TC_OPER_BYTES;
  static constexpr uint32_t BARS = STAGE0 + 2 * TC_STAGE_BYTES;  // 8 mbarriers
  static constexpr uint32_t TMEM_PTR = BARS + 8 * 8;
  static constexpr uint32_t RED = TMEM_PTR + 16;                 // [4 column parts][128 rows][2] floats
  static constexpr uint32_t LGX1 = RED + 4 * TC_M * 2 * 4;       // lgamma(x+1) table
  static constexpr uint32_t TOTAL = LGX1 + fm::LGX1_TABLE * 4;
};

// ---------------------------------------------------------------------------------------
// forward.  MODE 0: per-(split,row) online (max, sum-exp) of the scale logits.
//           MODE 1: per-(split,row) partial (-sum ll, S) given the merged (max, sum-exp).
// grid = (row tiles, gene splits); CTA handles gene tiles [t0, t1).
// ---------------------------------------------------------------------------------------
template <int MODE, bool ZINB>
__global__ void __launch_bounds__(TC_THREADS, 1)
    tc_decoder_fwd_kernel(const float* __restrict__ h, int64_t ldh, int64_t B, int64_t G,
                          const __half* __restrict__ packed, const float* __restrict__ gconst,
                          const float* __restrict__ x, int64_t ldx, const int64_t* __restrict__ row_idx,
                          const float* __restrict__ library, int tiles_per_split, int nsplit, float eps, int xvec,
                          float* __restrict__ lse_part /*[nsplit][B][2]*/,
                          float* __restrict__ out_part /*[nsplit][B][2]*/) {
  constexpr int NHEADS_PACK = ZINB ? 2 : 1;               // heads present in `packed`
  constexpr int NHEADS = (MODE == 1 && ZINB) ? 2 : 1;     // heads this kernel multiplies
  constexpr uint32_t TMEM_COLS = 2 * NHEADS * TC_N;       // 2 buffers
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((128u - (tc::smem_u32(smem_raw) & 127u)) & 127u);
  __half* s_hhi = reinterpret_cast<__half*>(smem + TcSmem::H_HI);
  __half* s_hlo = reinterpret_cast<__half*>(smem + TcSmem::H_LO);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + TcSmem::BARS);
  uint64_t *full = bars, *empty = bars + 2, *tfull = bars + 4, *tempty = bars + 6;
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(smem + TcSmem::TMEM_PTR);
  float* s_red = reinterpret_cast<float*>(smem + TcSmem::RED);
  float* s_lgx1 = reinterpret_cast<float*>(smem + TcSmem::LGX1);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int64_t r0 = (int64_t)blockIdx.x * TC_M;
  const int valid = (int)min((int64_t)TC_M, B - r0);
  const int nT = (int)((G + TC_N - 1) / TC_N);
  const int t0 = blockIdx.y * tiles_per_split;
  const int t1 = min(nT, t0 + tiles_per_split);
  const int ntiles = max(0, t1 - t0);
  const float* __restrict__ bias_pad = gconst + (int64_t)nT * TC_N * TC_GC;  // contiguous scale bias

  if (warp == 1 && lane == 0) {
    for (int i = 0; i < 2; ++i) {
      tc::mbar_init(&full[i], 1);
      tc::mbar_init(&empty[i], 1);
      tc::mbar_init(&tfull[i], 1);
      tc::mbar_init(&tempty[i], TC_EPI_WARPS);
    }
    tc::fence_mbar_init();
  }
  if (warp == 2) tc::tmem_alloc(s_tmem, TMEM_COLS);
  if (MODE == 1)
    for (int i = tid; i < fm::LGX1_TABLE; i += TC_THREADS) s_lgx1[i] = lgammaf((float)i + 1.f);
  tc_stage_h(h + r0 * ldh, ldh, valid, s_hhi, s_hlo, tid, TC_THREADS);
  tc::fence_proxy_async_smem();
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem_base = *s_tmem;

  if (warp == 0) {
    // ===== producer: one head tile (hi+lo, 64 KB) per stage
    if (lane == 0) {
      int it = 0;
      for (int t = t0; t < t1; ++t)
        for (int hd = 0; hd < NHEADS; ++hd, ++it) {
          const int s = it & 1;
          tc::mbar_wait(&empty[s], ((it >> 1) & 1) ^ 1);
          tc::mbar_arrive_expect_tx(&full[s], TC_STAGE_BYTES);
          const __half* src = packed + ((int64_t)(t * NHEADS_PACK + hd) * 2) * (TC_M * TC_H);
          uint8_t* dst = smem + TcSmem::STAGE0 + s * TC_STAGE_BYTES;
          tc::bulk_g2s(dst, src, TC_OPER_BYTES, &full[s]);
          tc::bulk_g2s(dst + TC_OPER_BYTES, src + TC_M * TC_H, TC_OPER_BYTES, &full[s]);
        }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: warp-uniform loop, one elected lane issues (tc.cuh)
    {
      constexpr uint32_t idesc = tc::idesc_f16_f32(TC_M, TC_N);
      const uint32_t leader = tc::elect_one();
      // A = h tile, B = weight tile of stage 0 (stage 1 = + TC_STAGE_BYTES): both K-major 128 x 128,
      // LBO = 128*16, SBO = 128, one MMA consumes K = 16 = 2 k-chunks
      const tc::Oper hA = tc::make_oper(tc::smem_u32(s_hhi), tc::smem_u32(s_hlo), TC_M * 16, 128, 2 * TC_M * 16);
      const uint32_t st0 = tc::smem_u32(smem + TcSmem::STAGE0);
      const tc::Oper wB = tc::make_oper(st0, st0 + TC_OPER_BYTES, TC_M * 16, 128, 2 * TC_M * 16);
      int it = 0;
      for (int j = 0; j < ntiles; ++j) {
        const int buf = j & 1;
        tc::mbar_wait(&tempty[buf], ((j >> 1) & 1) ^ 1);
        tc::fence_after_sync();
        for (int hd = 0; hd < NHEADS; ++hd, ++it) {
          const int s = it & 1;
          tc::mbar_wait(&full[s], (it >> 1) & 1);
          tc::fence_after_sync();
          if (leader) {
            tc::Oper b = wB;
            b.hi_lo += s * (TC_STAGE_BYTES >> 4);
            b.lo_lo += s * (TC_STAGE_BYTES >> 4);
            tc::mma3<TC_H / 16>(tmem_base + (buf * NHEADS + hd) * TC_N, hA, b, idesc, false);
            tc::mma_commit(&empty[s]);  // stage reusable once these MMAs have read it
          }
          __syncwarp();
        }
        if (leader) tc::mma_commit(&tfull[buf]);  // accumulators of tile j complete
        __syncwarp();
      }
    }
  } else if (warp >= 4) {
    // ===== epilogue: thread = (row r = 32*(warp%4) + lane, column part of 32 genes)
    const int q = warp & 3, part = (warp - 4) >> 2;
    const int r = 32 * q + lane;
    const bool rvalid = r < valid;
    const int64_t grow = rvalid ? (row_idx ? row_idx[r0 + r] : r0 + r) : 0;
    const float* xrow = x + grow * ldx;
    float m_run = -INFINITY, l_run = 0.f, nll = 0.f, S = 0.f;
    RowC rc;
    rc.m = 0.f; rc.invl = 0.f; rc.c = 0.f; rc.S = 0.f; rc.u = 0.f;
    float krow = 0.f;
    if (MODE == 1) {
      // merge the per-split (max, sum-exp) partials of this row
      float m = -INFINITY, l = 0.f;
      if (rvalid) {
        for (int sp = 0; sp < nsplit; ++sp) {
          const float2 ml = *reinterpret_cast<const float2*>(lse_part + ((int64_t)sp * B + r0 + r) * 2);
          if (ml.x == -INFINITY) continue;
          const float mn = fmaxf(m, ml.x);
          l = (m == -INFINITY ? 0.f : l * expf(m - mn)) + ml.y * expf(ml.x - mn);
          m = mn;
        }
        rc.m = m;
        rc.invl = 1.f / l;
        rc.c = expf(library[r0 + r]) / l;
        krow = (library[r0 + r] - m) * fm::LOG2E - log2f(l);
      }
    }
    const uint32_t lane_sel = (uint32_t)(32 * q) << 16;
    for (int j = 0; j < ntiles; ++j) {
      const int buf = j & 1;
      const int64_t g0 = (int64_t)(t0 + j) * TC_N;
      tc::mbar_wait(&tfull[buf], (j >> 1) & 1);
      tc::fence_after_sync();
#pragma unroll 1
      for (int c = 0; c < 2; ++c) {
        const int col0 = 32 * part + 16 * c;
        float sv[16], pv[16];
        const int64_t gb = g0 + col0;
        float4 b4[4];  // scale-head bias of this chunk (warp-uniform addresses): in flight behind the TMEM load
        if (MODE == 0) {
#pragma unroll
          for (int k = 0; k < 4; ++k) b4[k] = *reinterpret_cast<const float4*>(bias_pad + gb + 4 * k);
        }
        tc::tmem_ld16(tmem_base + lane_sel + (buf * NHEADS + 0) * TC_N + col0, sv);
        if (NHEADS == 2) tc::tmem_ld16(tmem_base + lane_sel + (buf * NHEADS + 1) * TC_N + col0, pv);
        tc::tmem_ld_wait();
        if (MODE == 0) {
          // log2 domain: t = logit * log2(e); running (max, sum 2^(t - max))
          float tm = -INFINITY;
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            sv[4 * k + 0] = fmaf(sv[4 * k + 0], TC_RESCALE * fm::LOG2E, b4[k].x * fm::LOG2E);
            sv[4 * k + 1] = fmaf(sv[4 * k + 1], TC_RESCALE * fm::LOG2E, b4[k].y * fm::LOG2E);
            sv[4 * k + 2] = fmaf(sv[4 * k + 2], TC_RESCALE * fm::LOG2E, b4[k].z * fm::LOG2E);
            sv[4 * k + 3] = fmaf(sv[4 * k + 3], TC_RESCALE * fm::LOG2E, b4[k].w * fm::LOG2E);
          }
          // genes of this chunk inside [0, G): 16 everywhere but at the ragged end (no per-element 64-bit test)
          const int nv = (int)max((int64_t)0, min((int64_t)16, G - gb));
          if (nv == 16) {
#pragma unroll
            for (int i = 0; i < 16; ++i) tm = fmaxf(tm, sv[i]);
          } else {
#pragma unroll
            for (int i = 0; i < 16; ++i)
              if (i < nv) tm = fmaxf(tm, sv[i]);
          }
          if (tm > m_run) {
            l_run *= fm::ex2(m_run - tm);  // m_run = -inf first time: 2^(-inf) = 0
            m_run = tm;
          }
          if (nv == 16) {
#pragma unroll
            for (int i = 0; i < 16; ++i) l_run += fm::ex2(sv[i] - m_run);
          } else {
#pragma unroll
            for (int i = 0; i < 16; ++i)
              if (i < nv) l_run += fm::ex2(sv[i] - m_run);
          }
        } else if (rvalid) {
          float xq[16];
          if (xvec) {  // 4 x 16-byte streaming loads (row start and G are multiples of 4 floats)
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              float4 t4 = make_float4(0.f, 0.f, 0.f, 0.f);
              if (gb + 4 * k < G) t4 = ld_stream4(xrow + gb + 4 * k);
              xq[4 * k] = t4.x; xq[4 * k + 1] = t4.y; xq[4 * k + 2] = t4.z; xq[4 * k + 3] = t4.w;
            }
          } else {
#pragma unroll
            for (int i = 0; i < 16; ++i) xq[i] = (gb + i < G) ? __ldg(xrow + gb + i) : 0.f;
          }
#pragma unroll 4
          for (int i = 0; i < 16; ++i) {
            if (gb + i >= G) break;
            const float* gcp = gconst + (gb + i) * TC_GC;
            const fm::GeneK gk = load_gene_k(gcp);
            const float s = fmaf(sv[i], TC_RESCALE, gcp[3]);
            const float pi = NHEADS == 2 ? fmaf(pv[i], TC_RESCALE, gcp[4]) : 0.f;
            float ll, g1, dpi, mu;
            fm::elem_fwd<ZINB>(xq[i], s, pi, krow, gk, eps, s_lgx1, ll, g1, dpi, mu);
            nll -= ll;
            S += g1;
          }
        }
      }
      tc::fence_before_sync();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(&tempty[buf]);
    }
    // combine the four column parts of each row and write this split's partial
    float a0 = MODE == 0 ? m_run * fm::LN2 : nll, a1 = MODE == 0 ? l_run : S;  // (max, sumexp) in natural units
    s_red[(part * TC_M + r) * 2 + 0] = a0;
    s_red[(part * TC_M + r) * 2 + 1] = a1;
    asm volatile("bar.sync 1, %0;" ::"n"(32 * TC_EPI_WARPS) : "memory");
    if (part == 0 && rvalid) {
      float2 o;
      if (MODE == 0) {
        float mn = a0;
#pragma unroll
        for (int p2 = 1; p2 < 4; ++p2) mn = fmaxf(mn, s_red[(p2 * TC_M + r) * 2]);
        float l = 0.f;
#pragma unroll
        for (int p2 = 0; p2 < 4; ++p2) {
          const float b0 = s_red[(p2 * TC_M + r) * 2], b1 = s_red[(p2 * TC_M + r) * 2 + 1];
          l += (b0 == -INFINITY) ? 0.f : b1 * expf(b0 - mn);
        }
        o.x = mn;
        o.y = l;
        *reinterpret_cast<float2*>(lse_part + ((int64_t)blockIdx.y * B + r0 + r) * 2) = o;
      } else {
        o.x = 0.f;
        o.y = 0.f;
#pragma unroll
        for (int p2 = 0; p2 < 4; ++p2) {
          o.x += s_red[(p2 * TC_M + r) * 2];
          o.y += s_red[(p2 * TC_M + r) * 2 + 1];
        }
        *reinterpret_cast<float2*>(out_part + ((int64_t)blockIdx.y * B + r0 + r) * 2) = o;
      }
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 2) tc::tmem_dealloc(tmem_base, TMEM_COLS);
}

// reconst[b] = sum_split nll; row_stats[b] = (max, sumexp, exp(library)/sumexp, sum_split S)
__global__ void tc_rowstats_combine_kernel(const float* __restrict__ lse_part, const float* __restrict__ out_part,
                                           const float* __restrict__ library, int64_t B, int nsplit,
                                           float* __restrict__ reconst, float* __restrict__ row_stats) {
  const int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  float m = -INFINITY, l = 0.f, nll = 0.f, S = 0.f;
  for (int sp = 0; sp < nsplit; ++sp) {
    const float2 ml = *reinterpret_cast<const float2*>(lse_part + ((int64_t)sp * B + b) * 2);
    const float2 o = *reinterpret_cast<const float2*>(out_part + ((int64_t)sp * B + b) * 2);
    nll += o.x;
    S += o.y;
    if (ml.x == -INFINITY) continue;
    const float mn = fmaxf(m, ml.x);
    l = (m == -INFINITY ? 0.f : l * expf(m - mn)) + ml.y * expf(ml.x - mn);
    m = mn;
  }
  reconst[b] = nll;
  *reinterpret_cast<float4*>(row_stats + b * 4) = make_float4(m, l, expf(library[b]) / l, S);
}

// ---------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------
struct TcPlan {
  int64_t nT, rowtiles;
  int nsplit, tiles_per_split;    // LSE / NLL-only kernels: 128-gene tiles
  int nsplit2, tiles_per_split2;  // fused NLL + dh kernel: 32-gene tiles
  size_t off_packed, off_gconst, off_lse, off_out, off_packed32, off_dh, off_fmax, off_aops, total;
};
size_t tc_aops_bytes(int64_t B, int64_t G, bool zinb);  // decoder_tc_train.cu

static TcPlan tc_plan(int64_t B, int64_t G, int nheads, bool aops = false) {
  TcPlan p;
  p.nT = cdiv(G, TC_N);
  p.rowtiles = cdiv(B, TC_M);
  int64_t ns = std::max<int64_t>(1, 148 / std::max<int64_t>(1, p.rowtiles));
  ns = std::min<int64_t>(ns, p.nT);
  p.tiles_per_split = (int)cdiv(p.nT, ns);
  p.nsplit = (int)cdiv(p.nT, p.tiles_per_split);
  auto al = [](size_t v) { return (v + 255) / 256 * 256; };
  p.off_packed = 0;
  p.off_gconst = al(p.off_packed + (size_t)p.nT * nheads * 2 * TC_M * TC_H * sizeof(__half));
  p.off_lse = al(p.off_gconst + (size_t)p.nT * TC_N * TC_GCONST_FLOATS * sizeof(float));
  const int64_t nT32 = cdiv(G, TC_NT);
  int64_t ns2 = std::min<int64_t>(std::max<int64_t>(1, 148 / std::max<int64_t>(1, p.rowtiles)), nT32);
  p.tiles_per_split2 = (int)cdiv(nT32, ns2);
  p.nsplit2 = (int)cdiv(nT32, p.tiles_per_split2);
  const int nsmax = std::max(p.nsplit, p.nsplit2);
  p.off_out = al(p.off_lse + (size_t)p.nsplit * B * 2 * sizeof(float));
  p.off_packed32 = al(p.off_out + (size_t)nsmax * B * 2 * sizeof(float));
  p.off_dh = al(p.off_packed32 + (size_t)p.nT * 4 * tc_w32_stage_bytes(nheads));
  // SCVI_PREC_AOPS: per-block maxima of the GEMM's row factors + the kept operands, behind everything else (the
  // offsets above do not depend on the flag)
  p.off_fmax = al(p.off_dh + (size_t)p.nsplit2 * 2 * B * TC_H * sizeof(float));
  p.off_aops = al(p.off_fmax + (size_t)cdiv(B, 8) * 3 * sizeof(float));
  p.total = aops ? al(p.off_aops + tc_aops_bytes(B, G, nheads == 2)) : p.off_fmax;
  return p;
}

// where the forward workspace keeps the packed 128-gene weight tiles and the gene-constant table
// (the backward reuses them when handed the forward workspace)
void tc_decoder_fwd_ws_offsets(int64_t B, int64_t G, int nheads, size_t* off_packed, size_t* off_gconst) {
  const TcPlan p = tc_plan(B, G, nheads);
  *off_packed = p.off_packed;
  *off_gconst = p.off_gconst;
}

int64_t tc_decoder_fwd_ws_bytes(int64_t B, int64_t H, int64_t G, int likelihood, bool aops) {
  (void)H;
  return (int64_t)tc_plan(B, G, likelihood == SCVI_LIK_ZINB ? 2 : 1, aops).total;
}
// where the forward workspace keeps the operands / factor maxima of SCVI_PREC_AOPS (read by the backward)
void tc_decoder_fwd_ws_aops(int64_t B, int64_t G, int nheads, size_t* off_fmax, size_t* off_aops) {
  const TcPlan p = tc_plan(B, G, nheads, true);
  *off_fmax = p.off_fmax;
  *off_aops = p.off_aops;
}

template <int MODE, bool ZINB>
static int tc_launch_fwd(const TcPlan& p, const float* h, int64_t ldh, int64_t B, int64_t G, const __half* packed,
                         const float* gconst, const float* x, int64_t ldx, const int64_t* row_idx,
                         const float* library, int xvec, float* lse, float* outp, cudaStream_t st) {
  auto k = tc_decoder_fwd_kernel<MODE, ZINB>;
  static std::atomic<uint64_t> optin{0};
  if (int rc = smem_optin((const void*)k, (int)TcSmem::TOTAL + 1024, optin, "tc_decoder_fwd")) return rc;
  dim3 grid((unsigned)p.rowtiles, (unsigned)p.nsplit);
  k<<<grid, TC_THREADS, TcSmem::TOTAL + 1024, st>>>(h, ldh, B, G, packed, gconst, x, ldx, row_idx, library,
                                                     p.tiles_per_split, p.nsplit, 1e-8f, xvec, lse, outp);
  return SCVI_OK;
}

int tc_launch_nll_dh(bool zinb, int xkind, const float* h, int64_t ldh, int64_t B, int64_t G, const uint8_t* packed32,
                     const void* x, int64_t ldx, const int64_t* row_idx, const float* library, const float* lse,
                     int nsplit_lse, int nsplit, int tiles_per_split, float* out_part, float* dh_part,
                     const float* g_reconst, float* reconst, float* row_stats, float* dh, int64_t lddh,
                     uint8_t* aops, float* fmax_part, cudaStream_t st);

// The parameter-only part of tc_decoder_nll_fwd (packed fp16 weight tiles, per-gene constants and tables)
// into the same workspace layout: lets the caller run it on another stream while the encoder computes.
int tc_decoder_nll_prepack(int64_t B, int64_t H, int64_t G, const float* Ws, const float* bs, const float* Wp,
                           const float* bp, const float* px_r, int likelihood, bool want_dh, void* ws, cudaStream_t st) {
  const bool zinb = likelihood == SCVI_LIK_ZINB;
  if (H != TC_H) {
    set_error("scvi_decoder_nll_prepack: n_hidden must be %d (got %lld)", TC_H, (long long)H);
    return SCVI_EUNSUPPORTED;
  }
  SCVI_REQUIRE(ws != nullptr && ((uintptr_t)ws % 256 == 0), "scvi_decoder_nll_prepack: ws must be 256-byte aligned");
  const int nheads = zinb ? 2 : 1;
  const TcPlan p = tc_plan(B, G, nheads);
  uint8_t* w8 = (uint8_t*)ws;
  const int64_t items = std::max<int64_t>(p.nT * nheads * TC_KC * TC_N, p.nT * TC_N * fm::LGT);
  tc_pack_heads_kernel<<<(unsigned)cdiv(items, 256), 256, 0, st>>>(Ws, bs, Wp, bp, px_r, G, nheads, 1e-8f,
                                                                    (__half*)(w8 + p.off_packed),
                                                                    want_dh ? w8 + p.off_packed32 : nullptr,
                                                                    (float*)(w8 + p.off_gconst));
  SCVI_LAUNCH_CHECK("scvi_decoder_nll_fwd(tc pack)");
  return SCVI_OK;
}

int tc_decoder_nll_fwd(const float* h, int64_t ldh, int64_t B, int64_t H, int64_t G, const float* Ws, const float* bs,
                       const float* Wp, const float* bp, const float* px_r, const void* xv, int64_t ldx,
                       const int64_t* row_idx, const float* library, int likelihood, float* reconst,
                       float* row_stats, const float* g_reconst, float* dh, int64_t lddh, void* ws,
                       cudaStream_t st, bool prepacked, int xkind) {
  const float* x = reinterpret_cast<const float*>(xv);  // the inference kernels below read fp32 counts only
  const bool keep_ops = (xkind & 16) != 0;  // SCVI_PREC_AOPS
  xkind &= 15;
  const bool zinb = likelihood == SCVI_LIK_ZINB;
  if (H != TC_H) {
    set_error("scvi_decoder_nll_fwd(SCVI_PREC_TC): n_hidden must be %d (got %lld)", TC_H, (long long)H);
    return SCVI_EUNSUPPORTED;
  }
  SCVI_REQUIRE(ws != nullptr && ((uintptr_t)ws % 256 == 0), "scvi_decoder_nll_fwd(SCVI_PREC_TC): ws must be 256-byte aligned");
  SCVI_REQUIRE(ldh % 4 == 0 && ((uintptr_t)h % 16 == 0), "scvi_decoder_nll_fwd(SCVI_PREC_TC): h must be 16-byte aligned");
  const int nheads = zinb ? 2 : 1;
  const TcPlan p = tc_plan(B, G, nheads);
  uint8_t* w8 = (uint8_t*)ws;
  __half* packed = (__half*)(w8 + p.off_packed);
  float* gconst = (float*)(w8 + p.off_gconst);
  float* lse = (float*)(w8 + p.off_lse);
  float* outp = (float*)(w8 + p.off_out);
  uint8_t* packed32 = dh ? w8 + p.off_packed32 : nullptr;
  if (!prepacked) {
    const int64_t items = std::max<int64_t>(p.nT * nheads * TC_KC * TC_N, p.nT * TC_N * fm::LGT);
    tc_pack_heads_kernel<<<(unsigned)cdiv(items, 256), 256, 0, st>>>(Ws, bs, Wp, bp, px_r, G, nheads, 1e-8f, packed, packed32, gconst);
    SCVI_LAUNCH_CHECK("scvi_decoder_nll_fwd(tc pack)");
  }
  const int xvec = (ldx % 4 == 0) && (G % 4 == 0) && ((uintptr_t)x % 16 == 0);
  int rc;
  if (dh) {  // training: LSE pass, then the fused NLL + dh kernel
    SCVI_REQUIRE(lddh >= TC_H && lddh % 4 == 0 && ((uintptr_t)dh % 16 == 0), "scvi_decoder_nll_fwd(SCVI_PREC_TC): dh must be 16-byte aligned");
    if (zinb) rc = tc_launch_fwd<0, true>(p, h, ldh, B, G, packed, gconst, x, ldx, row_idx, library, xvec, lse, outp, st);
    else rc = tc_launch_fwd<0, false>(p, h, ldh, B, G, packed, gconst, x, ldx, row_idx, library, xvec, lse, outp, st);
    if (rc) return rc;
    SCVI_LAUNCH_CHECK("scvi_decoder_nll_fwd(tc lse)");
    return tc_launch_nll_dh(zinb, xkind, h, ldh, B, G, packed32, xv, ldx, row_idx, library, lse, p.nsplit, p.nsplit2,
                            p.tiles_per_split2, outp, (float*)(w8 + p.off_dh), g_reconst, reconst, row_stats, dh,
                            lddh, keep_ops ? w8 + p.off_aops : nullptr, keep_ops ? (float*)(w8 + p.off_fmax) : nullptr, st);
  }
  if (xkind == 2) {
    set_error("scvi_decoder_nll_fwd(SCVI_X_U8): uint8 counts are read by the training kernels only (dh requested)");
    return SCVI_EUNSUPPORTED;
  }
  if (zinb) {
    if ((rc = tc_launch_fwd<0, true>(p, h, ldh, B, G, packed, gconst, x, ldx, row_idx, library, xvec, lse, outp, st))) return rc;
    SCVI_LAUNCH_CHECK("scvi_decoder_nll_fwd(tc lse)");
    if ((rc = tc_launch_fwd<1, true>(p, h, ldh, B, G, packed, gconst, x, ldx, row_idx, library, xvec, lse, outp, st))) return rc;
  } else {
    if ((rc = tc_launch_fwd<0, false>(p, h, ldh, B, G, packed, gconst, x, ldx, row_idx, library, xvec, lse, outp, st))) return rc;
    SCVI_LAUNCH_CHECK("scvi_decoder_nll_fwd(tc lse)");
    if ((rc = tc_launch_fwd<1, false>(p, h, ldh, B, G, packed, gconst, x, ldx, row_idx, library, xvec, lse, outp, st))) return rc;
  }
  SCVI_LAUNCH_CHECK("scvi_decoder_nll_fwd(tc nll)");
  tc_rowstats_combine_kernel<<<(unsigned)cdiv(B, 256), 256, 0, st>>>(lse, outp, library, B, p.nsplit, reconst, row_stats);
  SCVI_LAUNCH_CHECK("scvi_decoder_nll_fwd(tc combine)");
  return SCVI_OK;
}

}  // namespace scvi
