// tcgen05 / TMEM kernels of the encoder input layer (SCVI_LIN_TC):
//
//   E1  tc_enc_fwd_kernel   y[B,128] = log1p(x) We^T   (CTA = 128 rows x a range of 32-gene K-tiles)
//       16 converter warps read the minibatch rows straight from the resident count matrix, apply log1p,
//       split to fp16 hi/lo and stage the A operand in shared memory; We K-tiles arrive pre-packed by bulk
//       copy; the accumulator stays in TMEM over the whole K loop.  Row sums (library) on the side.
//   E2  tc_enc_dw_kernel    dWe[128,G] = dy^T log1p(x)  (CTA = 128 genes x a range of 32-row K-tiles)
//       the same converter pattern stages log1p(x)^T as an MN-major B operand; dy tiles arrive pre-packed
//       and are used as an MN-major A operand.
//
// Replace nn.Linear of the encoder's FCLayers layer 0 and the pre-ops log1p / library
// (reference src/scvi/nn/_base_components.py:148-188, src/scvi/module/_vae.py:370-374) and their autograd
// backward.  fp32-grade accuracy through the 3-MMA fp16 split (tc_layout.cuh).
#include "tc_layout.cuh"

namespace scvi {

constexpr int EN_CONV_WARPS = 16;
// raw-count loaders: warps 0, 2 and the two warps after the converters.  Each warp instruction copies one
// 512-byte row segment with 32 x cp.async 16 B (LDGSTS): tools/ubench_gather.cu measures 4.5 TB/s chip-wide
// for this scheme against 2.4 TB/s for one 512-byte bulk copy per row (the TMA engine retires ~32 M copies/s/SM)
constexpr int EN_LOAD_WARPS = 4;
constexpr int EN_THREADS = 32 * (4 + EN_CONV_WARPS + 2);
__device__ __forceinline__ int en_loader_id(int warp) {
  return warp == 0 ? 0 : warp == 2 ? 1 : warp >= 4 + EN_CONV_WARPS ? 2 + warp - (4 + EN_CONV_WARPS) : -1;
}
constexpr int EN_NS = 4;                          // pipeline stages
constexpr uint32_t EN_BLK = TC_M * TC_NT * 2;     // 128 x 32 fp16 block = 8 KB
constexpr int EN_X_SCALE_LOG2 = 11;               // log1p(x) <= 16 -> < 2^15

// ---------------------------------------------------------------------------------------
// pack We fp32 [128, ldw] -> per 32-gene K-tile {hi, lo} blocks, addr(n, kc) = kc * 2048 + n * 16
// ---------------------------------------------------------------------------------------
__global__ void tc_pack_we_kernel(const float* __restrict__ W, int64_t ldw, int64_t K, int64_t nKT,
                                  uint8_t* __restrict__ out) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= nKT * 4 * TC_M) return;
  const int n = (int)(e % TC_M);
  const int kc = (int)((e / TC_M) % 4);
  const int64_t t = e / (4 * TC_M);
  float v[8];
#pragma unroll
  for (int q = 0; q < 8; ++q) {
    const int64_t k = t * TC_NT + 8 * kc + q;
    v[q] = (k < K) ? fminf(fmaxf(W[n * ldw + k] * TC_SCALE_W, -60000.f), 60000.f) : 0.f;
  }
  uint4 hi, lo;
  tc::split8(v, hi, lo);
  uint8_t* base = out + t * 2 * EN_BLK;
  *reinterpret_cast<uint4*>(base + kc * (TC_M * 16) + n * 16) = hi;
  *reinterpret_cast<uint4*>(base + EN_BLK + kc * (TC_M * 16) + n * 16) = lo;
}

// raw fp32 count tiles, gathered row by row with 1-D bulk copies (TMA engine): rows padded by 16 B so the
// converters' 16-byte shared loads are bank-conflict free
constexpr uint32_t EN_RAW1_ROW = TC_NT * 4 + 16;          // E1: 128 rows x 32 genes
constexpr uint32_t EN_RAW1 = TC_M * EN_RAW1_ROW;          // 18 KB
constexpr uint32_t EN_RAW2_ROW = TC_M * 4 + 16;           // E2: 32 rows x 128 genes
constexpr uint32_t EN_RAW2 = TC_NT * EN_RAW2_ROW;         // 16.5 KB
struct EnSmem {
  static constexpr uint32_t A = 0;                       // EN_NS x {hi, lo}
  static constexpr uint32_t W = EN_NS * 2 * EN_BLK;      // EN_NS x {hi, lo}
  static constexpr uint32_t RAW = 2 * EN_NS * 2 * EN_BLK; // EN_NS raw tiles
  static constexpr uint32_t BARS = RAW + EN_NS * EN_RAW1;
  static constexpr int NRAW8 = 12;                       // uint8 raw ring: 12 x 4.5 KB tiles
  static constexpr uint32_t TMEM_PTR = BARS + 48 * 8;
  static constexpr uint32_t RED = TMEM_PTR + 16;         // [4][128] row-sum partials
  static constexpr uint32_t LUT = RED + 4 * TC_M * 4;    // 256 x (hi | lo << 16), uint8 counts
  static constexpr uint32_t TOTAL = LUT + 256 * 4;
};
static_assert(EN_RAW2 <= EN_RAW1, "raw stage size");

// load 8 consecutive counts of one row (zero beyond G)
__device__ __forceinline__ void en_load8(const float* __restrict__ xrow, int64_t gb, int64_t G, bool rvalid, int xvec,
                                         float (&v)[8]) {
  if (rvalid && xvec) {
    const float4 a = (gb < G) ? ld_stream4(xrow + gb) : make_float4(0.f, 0.f, 0.f, 0.f);
    const float4 b = (gb + 4 < G) ? ld_stream4(xrow + gb + 4) : make_float4(0.f, 0.f, 0.f, 0.f);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  } else {
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = (rvalid && gb + i < G) ? __ldg(xrow + gb + i) : 0.f;
  }
}

// ---- uint8 counts (SCVI_LIN_X_U8): the loaders move 1 byte per count (128-byte row segments, 8 x cp.async 16 B,
// four rows per warp instruction) and the converters replace log1p + scale + clamp + fp16 split by ONE shared-memory
// look-up per count: lut[c] = (hi | lo << 16) of exactly the value the fp32 path computes for the count c (built with
// the same flog / split8 at kernel start: the MMA operands, hence the results, are bit-identical to the fp32 path).
constexpr uint32_t EN_RAW8_ROW = TC_M + 16;  // 128 counts + 16 B pad (cp.async needs 16-byte aligned rows)
__device__ __forceinline__ void en_build_lut(uint32_t* lut, int log1p_in, int tid, int nthreads) {
  const float sX = tc::pow2i(EN_X_SCALE_LOG2);
  for (int t = tid; t < 32; t += nthreads) {
    float v[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float xc = (float)(8 * t + i);
      v[i] = (log1p_in ? fm::flog(1.f + xc) : xc) * sX;
      v[i] = fminf(fmaxf(v[i], -60000.f), 60000.f);
    }
    uint4 hi, lo;
    tc::split8(v, hi, lo);
    const uint32_t h[4] = {hi.x, hi.y, hi.z, hi.w}, l[4] = {lo.x, lo.y, lo.z, lo.w};
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      lut[8 * t + 2 * p] = __byte_perm(h[p], l[p], 0x5410);
      lut[8 * t + 2 * p + 1] = __byte_perm(h[p], l[p], 0x7632);
    }
  }
}
// 8 counts (two 32-bit words of bytes) -> the staged operand chunk {8 hi halves, 8 lo halves}
__device__ __forceinline__ void en_lut8(const uint32_t* __restrict__ lut, uint32_t b0, uint32_t b1, uint4& hi, uint4& lo) {
  const uint32_t w0 = lut[b0 & 0xff], w1 = lut[(b0 >> 8) & 0xff], w2 = lut[(b0 >> 16) & 0xff], w3 = lut[b0 >> 24];
  const uint32_t w4 = lut[b1 & 0xff], w5 = lut[(b1 >> 8) & 0xff], w6 = lut[(b1 >> 16) & 0xff], w7 = lut[b1 >> 24];
  hi = make_uint4(__byte_perm(w0, w1, 0x5410), __byte_perm(w2, w3, 0x5410), __byte_perm(w4, w5, 0x5410), __byte_perm(w6, w7, 0x5410));
  lo = make_uint4(__byte_perm(w0, w1, 0x7632), __byte_perm(w2, w3, 0x7632), __byte_perm(w4, w5, 0x7632), __byte_perm(w6, w7, 0x7632));
}
// keep the first n (0..8) bytes of the pair (b0, b1)
__device__ __forceinline__ void en_keep_bytes(uint32_t& b0, uint32_t& b1, int n) {
  n = max(0, min(8, n));
  b0 &= n >= 4 ? 0xffffffffu : ((1u << (8 * n)) - 1u);
  b1 &= n >= 8 ? 0xffffffffu : (n <= 4 ? 0u : ((1u << (8 * (n - 4))) - 1u));
}

// ---------------------------------------------------------------------------------------
// E1: forward
// ---------------------------------------------------------------------------------------
// E1 shared memory: 2 stages of {A hi, A lo} and {We hi, We lo} (32-gene K-tiles) + 2 raw chunks of
// 128 rows x 128 genes fp32 (one 512-byte bulk copy per row: the TMA engine is issue-bound on tiny copies)
struct En1Smem {
  static constexpr int NS = 2;
  static constexpr uint32_t A = 0;
  static constexpr uint32_t W = NS * 2 * EN_BLK;
  static constexpr uint32_t RAW = 2 * NS * 2 * EN_BLK;
  static constexpr uint32_t RAW_ROW = TC_M * 4 + 16;
  static constexpr uint32_t RAW_BYTES = TC_M * RAW_ROW;  // 66 KB
  static constexpr uint32_t BARS = RAW + 2 * RAW_BYTES;
  static constexpr int NRAW8 = 2;
  static constexpr uint32_t TMEM_PTR = BARS + 32 * 8;
  static constexpr uint32_t RED = TMEM_PTR + 16;
  static constexpr uint32_t LUT = RED + 4 * TC_M * 4;    // 256 x (hi | lo << 16)
  static constexpr uint32_t RAW8_BYTES = TC_M * EN_RAW8_ROW;  // uint8 chunk: 18 KB
  static constexpr uint32_t TOTAL = LUT + 256 * 4;
};

// uint8 counts: the raw chunks are 4x smaller, which buys 4 operand stages (the packed We tiles come from L2 with
// ~1 us of bulk-copy latency each: two stages in flight kept the tensor pipe waiting) and a 4-deep raw ring
struct En1SmemU8 {
  static constexpr int NS = 4;
  static constexpr uint32_t A = 0;
  static constexpr uint32_t W = NS * 2 * EN_BLK;
  static constexpr uint32_t RAW = 2 * NS * 2 * EN_BLK;
  static constexpr int NRAW8 = 4;
  static constexpr uint32_t RAW8_BYTES = TC_M * EN_RAW8_ROW;
  static constexpr uint32_t RAW_BYTES = RAW8_BYTES, RAW_ROW = EN_RAW8_ROW;
  static constexpr uint32_t BARS = RAW + NRAW8 * RAW8_BYTES;
  static constexpr uint32_t TMEM_PTR = BARS + 32 * 8;
  static constexpr uint32_t RED = TMEM_PTR + 16;
  static constexpr uint32_t LUT = RED + 4 * TC_M * 4;
  static constexpr uint32_t TOTAL = LUT + 256 * 4;
};
static_assert(En1SmemU8::TOTAL + 128 <= 232448, "E1 uint8 shared memory");
template <bool U8>
struct En1Layout { using type = En1Smem; };
template <>
struct En1Layout<true> { using type = En1SmemU8; };

template <bool U8>
__global__ void __launch_bounds__(EN_THREADS, 1)
    tc_enc_fwd_kernel(const void* __restrict__ xv, int64_t ldx, const int64_t* __restrict__ row_idx, int64_t B,
                      int64_t G, const uint8_t* __restrict__ wpack, int chunks_per_split, int log1p_in,
                      float* __restrict__ ypart /*[nsplit][B][128]*/, float* __restrict__ rowsum_part /*[nsplit][B]*/) {
  using L = typename En1Layout<U8>::type;
  constexpr int NS = L::NS;  // operand stages
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((128u - (tc::smem_u32(smem_raw) & 127u)) & 127u);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::BARS);
  uint64_t *wfull = bars, *afull = bars + 4, *empty = bars + 8, *accfull = bars + 12, *rfull = bars + 16, *rempty = bars + 24;
  constexpr int NRAW = U8 ? L::NRAW8 : 2;  // raw-count ring depth
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(smem + L::TMEM_PTR);
  float* s_red = reinterpret_cast<float*>(smem + L::RED);
  uint32_t* s_lut = reinterpret_cast<uint32_t*>(smem + L::LUT);
  const float* x = reinterpret_cast<const float*>(xv);
  const uint8_t* x8 = reinterpret_cast<const uint8_t*>(xv);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int64_t r0 = (int64_t)blockIdx.x * TC_M;
  const int valid = (int)min((int64_t)TC_M, B - r0);
  const int nKC = (int)((G + TC_M - 1) / TC_M);  // 128-gene chunks
  const int c0 = blockIdx.y * chunks_per_split;
  const int nchunks = max(0, min(nKC, c0 + chunks_per_split) - c0);
  const int ntiles = 4 * nchunks;  // 32-gene K-tiles (the packed We covers whole chunks, zero padded)

  if (warp == 1 && lane == 0) {
    for (int i = 0; i < NS; ++i) {
      tc::mbar_init(&wfull[i], 1);
      tc::mbar_init(&afull[i], U8 ? 4 : EN_CONV_WARPS);  // uint8: one group of 4 warps converts a K-tile
      tc::mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < NRAW; ++i) {
      tc::mbar_init(&rfull[i], 32 * EN_LOAD_WARPS);
      tc::mbar_init(&rempty[i], EN_CONV_WARPS);
    }
    tc::mbar_init(accfull, 1);
    tc::fence_mbar_init();
  }
  if (warp == 2) tc::tmem_alloc(s_tmem, 128);
  if (U8) en_build_lut(s_lut, log1p_in, tid, EN_THREADS);
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem_base = *s_tmem;
  const int lid = en_loader_id(warp);

  if (lid >= 0 && U8) {
    // ===== raw producers, uint8 counts: a row segment of a chunk is 128 B = 8 x 16 B, one warp instruction copies
    // the segments of 4 rows (lane = (row lane / 8, piece lane % 8)); lane owns the pointers of rows lane + 32 k
    const uint8_t* rp[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int rr = lane + 32 * k;
      rp[k] = (rr < valid) ? x8 + (row_idx ? row_idx[r0 + rr] : r0 + rr) * ldx : nullptr;
    }
    const int piece = lane & 7, rsub = lane >> 3;
    for (int c = 0; c < nchunks; ++c) {
      const int s = c % NRAW;
      const int64_t gt = (int64_t)(c0 + c) * TC_M;
      const uint32_t rowbytes = (uint32_t)min((int64_t)TC_M, G - gt);
      const bool lane_on = (uint32_t)piece * 16 < rowbytes;
      tc::mbar_wait(&rempty[s], ((c / NRAW) & 1) ^ 1);
      const uint32_t dst0 = tc::smem_u32(smem + L::RAW + s * L::RAW8_BYTES) + piece * 16;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
#pragma unroll
        for (int i = 0; i < 8 / EN_LOAD_WARPS; ++i) {
          const int rr = 4 * (lid + EN_LOAD_WARPS * i) + rsub;  // row 32 k + rr of the tile
          const uint8_t* p = reinterpret_cast<const uint8_t*>(
              __shfl_sync(0xffffffffu, reinterpret_cast<unsigned long long>(rp[k]), rr));
          if (p && lane_on) tc::cp_async16(dst0 + (32 * k + rr) * EN_RAW8_ROW, p + gt + piece * 16);
        }
      }
      tc::cp_async_mbar_arrive_noinc(&rfull[s]);
    }
  } else if (lid >= 0) {
    // ===== raw producers: per chunk, loader warp `lid` copies the 512-byte segments of rows lid, lid + 4, ...
    // (one warp instruction = one row segment; lane owns the row pointers of rows lane + 32 k)
    const float* rp[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int rr = lane + 32 * k;
      rp[k] = (rr < valid) ? x + (row_idx ? row_idx[r0 + rr] : r0 + rr) * ldx : nullptr;
    }
    for (int c = 0; c < nchunks; ++c) {
      const int s = c & 1;
      const int64_t gt = (int64_t)(c0 + c) * TC_M;
      const uint32_t rowbytes = (uint32_t)min((int64_t)TC_M, G - gt) * 4;
      const bool lane_on = (uint32_t)lane * 16 < rowbytes;
      tc::mbar_wait(&rempty[s], ((c >> 1) & 1) ^ 1);
      const uint32_t dst0 = tc::smem_u32(smem + L::RAW + s * L::RAW_BYTES) + lane * 16;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
#pragma unroll
        for (int i = 0; i < 32 / EN_LOAD_WARPS; ++i) {
          const int sl = lid + EN_LOAD_WARPS * i;
          const float* p = reinterpret_cast<const float*>(
              __shfl_sync(0xffffffffu, reinterpret_cast<unsigned long long>(rp[k]), sl));
          if (p && lane_on) tc::cp_async16(dst0 + (32 * k + sl) * L::RAW_ROW, p + gt + lane * 4);
        }
      }
      tc::cp_async_mbar_arrive_noinc(&rfull[s]);
    }
  } else if (warp == 3) {
    // ===== We producer: one packed 32-gene K-tile per stage
    if (lane == 0) {
      for (int j = 0; j < ntiles; ++j) {
        const int s = j % NS;
        tc::mbar_wait(&empty[s], ((j / NS) & 1) ^ 1);
        tc::mbar_arrive_expect_tx(&wfull[s], 2 * EN_BLK);
        tc::bulk_g2s(smem + L::W + s * 2 * EN_BLK, wpack + ((size_t)c0 * 4 + j) * 2 * EN_BLK, 2 * EN_BLK, &wfull[s]);
      }
    }
  } else if (warp == 1) {
    if (ntiles > 0) {  // warp-uniform issue loop, one elected lane issues (tc.cuh)
      constexpr uint32_t idesc = tc::idesc_f16_f32(TC_M, TC_H, 0, 0);
      const uint32_t leader = tc::elect_one();
      // stage 0 operands (stage 1 = + 2 * EN_BLK): A = converted log1p(x) tile, B = We K-tile; both K-major
      // 128 x 32: LBO = 128*16, SBO = 128, K-step = 2 chunks
      const uint32_t a0 = tc::smem_u32(smem + L::A), w0 = tc::smem_u32(smem + L::W);
      const tc::Oper A0 = tc::make_oper(a0, a0 + EN_BLK, TC_M * 16, 128, 2 * TC_M * 16);
      const tc::Oper W0 = tc::make_oper(w0, w0 + EN_BLK, TC_M * 16, 128, 2 * TC_M * 16);
      for (int j = 0; j < ntiles; ++j) {
        const int s = j % NS;
        tc::mbar_wait(&wfull[s], (j / NS) & 1);
        tc::mbar_wait(&afull[s], (j / NS) & 1);
        tc::fence_after_sync();
        if (leader) {
          tc::Oper a = A0, w = W0;
          a.hi_lo += s * ((2 * EN_BLK) >> 4); a.lo_lo += s * ((2 * EN_BLK) >> 4);
          w.hi_lo += s * ((2 * EN_BLK) >> 4); w.lo_lo += s * ((2 * EN_BLK) >> 4);
          tc::mma3<TC_NT / 16>(tmem_base, a, w, idesc, j > 0);
          tc::mma_commit(&empty[s]);
        }
        __syncwarp();
      }
      if (leader) tc::mma_commit(accfull);
      __syncwarp();
    }
  } else if (warp >= 4) {
    // ===== converters: thread = (row r, chunk cg of 8 genes inside each 32-gene K-tile)
    const int q = warp & 3, cg = (warp - 4) >> 2;
    const int r = 32 * q + lane;
    const bool rvalid = r < valid;
    const float sX = tc::pow2i(EN_X_SCALE_LOG2);
    float rowsum = 0.f;
    uint32_t isum = 0;  // uint8 counts: exact integer row sum (< 2^24: equals the fp32 sum bit for bit)
    // uint8: converter group cg (4 warps = the 128 rows) takes K-tile `cg` of every 128-gene chunk: a thread converts
    // the 32 counts of its row in that tile and pays one proxy fence + one barrier hand-off for them
    for (int c = 0; U8 && c < nchunks; ++c) {
      const int rs = c % NRAW;
      tc::mbar_wait(&rfull[rs], (c / NRAW) & 1);
      const uint8_t* raw = smem + L::RAW + rs * L::RAW8_BYTES + r * EN_RAW8_ROW + cg * TC_NT;
      const int64_t gt = (int64_t)(c0 + c) * TC_M + cg * TC_NT;  // first gene of this thread's 32
      const int j = 4 * c + cg, s = j % NS;
      uint4 b0 = *reinterpret_cast<const uint4*>(raw);
      uint4 b1 = *reinterpret_cast<const uint4*>(raw + 16);
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(&rempty[rs]);  // raw chunk consumed by this warp
      if (gt + TC_NT > G) {  // genes beyond G were not copied: stale bytes
        en_keep_bytes(b0.x, b0.y, (int)(G - gt));
        en_keep_bytes(b0.z, b0.w, (int)(G - gt - 8));
        en_keep_bytes(b1.x, b1.y, (int)(G - gt - 16));
        en_keep_bytes(b1.z, b1.w, (int)(G - gt - 24));
      }
      isum = __dp4a(b0.x, 0x01010101u, isum); isum = __dp4a(b0.y, 0x01010101u, isum);
      isum = __dp4a(b0.z, 0x01010101u, isum); isum = __dp4a(b0.w, 0x01010101u, isum);
      isum = __dp4a(b1.x, 0x01010101u, isum); isum = __dp4a(b1.y, 0x01010101u, isum);
      isum = __dp4a(b1.z, 0x01010101u, isum); isum = __dp4a(b1.w, 0x01010101u, isum);
      uint4 hi[4], lo[4];
      en_lut8(s_lut, b0.x, b0.y, hi[0], lo[0]);
      en_lut8(s_lut, b0.z, b0.w, hi[1], lo[1]);
      en_lut8(s_lut, b1.x, b1.y, hi[2], lo[2]);
      en_lut8(s_lut, b1.z, b1.w, hi[3], lo[3]);
      tc::mbar_wait(&empty[s], ((j / NS) & 1) ^ 1);
      uint8_t* a = smem + L::A + s * 2 * EN_BLK + r * 16;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        *reinterpret_cast<uint4*>(a + k * (TC_M * 16)) = hi[k];
        *reinterpret_cast<uint4*>(a + k * (TC_M * 16) + EN_BLK) = lo[k];
      }
      tc::fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(&afull[s]);
    }
    if (U8) rowsum = (float)isum;
    for (int c = 0; !U8 && c < nchunks; ++c) {
      const int rs = c & 1;
      tc::mbar_wait(&rfull[rs], (c >> 1) & 1);
      const uint8_t* raw = smem + L::RAW + rs * L::RAW_BYTES + r * L::RAW_ROW + cg * 32;
#pragma unroll
      for (int sub = 0; sub < 4; ++sub) {
        const int j = 4 * c + sub, s = j % NS;
        const int64_t gb = (int64_t)(c0 + c) * TC_M + sub * TC_NT + 8 * cg;
        const float4 x0 = *reinterpret_cast<const float4*>(raw + sub * 128);
        const float4 x1 = *reinterpret_cast<const float4*>(raw + sub * 128 + 16);
        float xc[8] = {x0.x, x0.y, x0.z, x0.w, x1.x, x1.y, x1.z, x1.w};
        float v[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          xc[i] = (rvalid && gb + i < G) ? xc[i] : 0.f;  // rows / genes beyond the edge were not copied
          rowsum += xc[i];
          v[i] = (log1p_in ? fm::flog(1.f + xc[i]) : xc[i]) * sX;
          v[i] = fminf(fmaxf(v[i], -60000.f), 60000.f);
        }
        uint4 hi, lo;
        tc::split8(v, hi, lo);
        tc::mbar_wait(&empty[s], ((j / NS) & 1) ^ 1);
        uint8_t* a = smem + L::A + s * 2 * EN_BLK + cg * (TC_M * 16) + r * 16;
        *reinterpret_cast<uint4*>(a) = hi;
        *reinterpret_cast<uint4*>(a + EN_BLK) = lo;
        tc::fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) tc::mbar_arrive(&afull[s]);
      }
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(&rempty[rs]);  // raw chunk consumed by this warp
    }
    s_red[cg * TC_M + r] = rowsum;
    asm volatile("bar.sync 1, %0;" ::"n"(32 * EN_CONV_WARPS) : "memory");
    if (cg == 0 && rvalid)
      rowsum_part[(int64_t)blockIdx.y * B + r0 + r] = (s_red[r] + s_red[TC_M + r]) + (s_red[2 * TC_M + r] + s_red[3 * TC_M + r]);
    if (ntiles > 0) {
      tc::mbar_wait(accfull, 0);
      tc::fence_after_sync();
    }
    const uint32_t lane_sel = (uint32_t)(32 * q) << 16;
    const float k = tc::pow2i(-EN_X_SCALE_LOG2) * (1.f / TC_SCALE_W);
    float* o = ypart + ((int64_t)blockIdx.y * B + r0 + r) * TC_H + 32 * cg;
#pragma unroll 1
    for (int c = 0; c < 4; ++c) {
      float v[8];
      if (ntiles > 0) {
        tc::tmem_ld8(tmem_base + lane_sel + 32 * cg + 8 * c, v);
        tc::tmem_ld_wait();
      } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] = 0.f;
      }
      if (rvalid) {
        *reinterpret_cast<float4*>(o + 8 * c) = make_float4(v[0] * k, v[1] * k, v[2] * k, v[3] * k);
        *reinterpret_cast<float4*>(o + 8 * c + 4) = make_float4(v[4] * k, v[5] * k, v[6] * k, v[7] * k);
      }
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 2) tc::tmem_dealloc(tmem_base, 128);
}

// y = sum_split ypart + bias; library = log(sum_split rowsum); per-64-row-tile BatchNorm statistics.
// block = one BN tile (64 rows) x 128 columns, 512 threads: thread (col, row group) owns 16 rows of one column.
__global__ void __launch_bounds__(512)
    tc_enc_fwd_combine_kernel(const float* __restrict__ ypart, const float* __restrict__ rowsum_part, int nsplit,
                              int64_t B, const float* __restrict__ bias, float* __restrict__ y, int64_t ldy,
                              float* __restrict__ library, float* __restrict__ bn_partials) {
  __shared__ float s_scr[4][TC_H];
  const int col = threadIdx.x & 127, rg = threadIdx.x >> 7;
  const int64_t r0 = (int64_t)blockIdx.x * SCVI_BN_TILE;
  const int cnt = (int)min((int64_t)SCVI_BN_TILE, B - r0);
  const float bv = bias ? bias[col] : 0.f;
  // split outermost, the thread's 16 rows innermost: 16 independent loads in flight per split (the per-element
  // summation order bias + split 0 + split 1 + ... is unchanged)
  float vals[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) vals[i] = (rg * 16 + i < cnt) ? bv : 0.f;
  for (int sp = 0; sp < nsplit; ++sp) {
    const float* yp = ypart + ((int64_t)sp * B + r0 + rg * 16) * TC_H + col;
#pragma unroll
    for (int i = 0; i < 16; ++i)
      if (rg * 16 + i < cnt) vals[i] += yp[i * TC_H];
  }
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 16; ++i)
    if (rg * 16 + i < cnt) {
      y[(r0 + rg * 16 + i) * ldy + col] = vals[i];
      s += vals[i];
    }
  s_scr[rg][col] = s;
  __syncthreads();
  const float mean = ((s_scr[0][col] + s_scr[1][col]) + (s_scr[2][col] + s_scr[3][col])) / (float)cnt;
  __syncthreads();
  float m2 = 0.f;
#pragma unroll
  for (int i = 0; i < 16; ++i)
    if (rg * 16 + i < cnt) {
      const float d = vals[i] - mean;
      m2 = fmaf(d, d, m2);
    }
  s_scr[rg][col] = m2;
  __syncthreads();
  if (rg == 0 && bn_partials) {
    bn_partials[((int64_t)blockIdx.x * 2 + 0) * TC_H + col] = mean;
    bn_partials[((int64_t)blockIdx.x * 2 + 1) * TC_H + col] = (s_scr[0][col] + s_scr[1][col]) + (s_scr[2][col] + s_scr[3][col]);
  }
  if (library && threadIdx.x < cnt) {
    float t = 0.f;
#pragma unroll 8
    for (int sp = 0; sp < nsplit; ++sp) t += rowsum_part[(int64_t)sp * B + r0 + threadIdx.x];
    library[r0 + threadIdx.x] = logf(t);
  }
}

// ---------------------------------------------------------------------------------------
// E2: dWe = dy^T log1p(x)
// ---------------------------------------------------------------------------------------
// max |src[b, 0..N)| per block -> part[blockIdx.x] (no atomics, no zero-fill).  N % 4 == 0, rows 16-byte aligned.
__global__ void __launch_bounds__(256)
    tc_absmax_kernel(const float* __restrict__ src, int64_t ld, int64_t B, int N, float* __restrict__ part) {
  __shared__ float s_m[8];
  float m = 0.f;
  const int n4 = N >> 2;
  const int64_t total = B * n4;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    const float4 v = *reinterpret_cast<const float4*>(src + (e / n4) * ld + 4 * (e % n4));
    m = fmaxf(fmaxf(m, fmaxf(fabsf(v.x), fabsf(v.y))), fmaxf(fabsf(v.z), fabsf(v.w)));
  }
  m = warp_max(m);
  if ((threadIdx.x & 31) == 0) s_m[threadIdx.x >> 5] = m;
  __syncthreads();
  if (threadIdx.x < 32) {
    m = threadIdx.x < 8 ? s_m[threadIdx.x] : 0.f;
    m = warp_max(m);
    if (threadIdx.x == 0) part[blockIdx.x] = m;
  }
}

// dy fp32 [B, 128] -> per 32-row tile {hi, lo} blocks, addr(r, kc) = kc * 512 + r * 16, scaled by
// 2^(14 - ceil(log2(absmax))), absmax = max over the n_part partial maxima (also written to *absmax_out)
__global__ void __launch_bounds__(512)
    tc_pack_rows32_kernel(const float* __restrict__ src, int64_t ld, int64_t B, const float* __restrict__ absmax_part,
                          int n_part, float* __restrict__ absmax_out, uint8_t* __restrict__ out) {
  __shared__ float s_am;
  const int64_t t = blockIdx.x;
  const int r = threadIdx.x & 31, kc = threadIdx.x >> 5;
  const int64_t b = t * TC_NT + r;
  if (kc == 0) {
    float m = 0.f;
    for (int i = r; i < n_part; i += 32) m = fmaxf(m, absmax_part[i]);
    m = warp_max(m);
    if (r == 0) {
      s_am = m;
      if (t == 0) *absmax_out = m;
    }
  }
  __syncthreads();
  const int e = max(-100, min(100, (int)ceilf(log2f(fmaxf(s_am, 1e-30f)))));
  const float sc = tc::pow2i(14 - e);
  float v[8];
  if (b < B) {
    const float4 a = *reinterpret_cast<const float4*>(src + b * ld + 8 * kc);
    const float4 c = *reinterpret_cast<const float4*>(src + b * ld + 8 * kc + 4);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = c.x; v[5] = c.y; v[6] = c.z; v[7] = c.w;
#pragma unroll
    for (int q = 0; q < 8; ++q) v[q] *= sc;
  } else {
#pragma unroll
    for (int q = 0; q < 8; ++q) v[q] = 0.f;
  }
  uint4 hi, lo;
  tc::split8(v, hi, lo);
  uint8_t* base = out + t * 2 * TC_BLK32;
  *reinterpret_cast<uint4*>(base + kc * TC_NT * 16 + r * 16) = hi;
  *reinterpret_cast<uint4*>(base + TC_BLK32 + kc * TC_NT * 16 + r * 16) = lo;
}

template <bool U8>
__global__ void __launch_bounds__(EN_THREADS, 1)
    tc_enc_dw_kernel(const void* __restrict__ xv, int64_t ldx, const int64_t* __restrict__ row_idx, int64_t B,
                     int64_t G, const uint8_t* __restrict__ dypack, const float* __restrict__ absmax,
                     int tiles_per_split, int log1p_in, int xvec, float* __restrict__ slab /*[nsplit][128][G]*/) {
  using L = EnSmem;  // A region = dy tiles (bulk copied), W region = converted log1p(x)^T tiles
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((128u - (tc::smem_u32(smem_raw) & 127u)) & 127u);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::BARS);
  uint64_t *dfull = bars, *xfull = bars + EN_NS, *empty = bars + 2 * EN_NS, *accfull = bars + 3 * EN_NS;
  uint64_t *rfull = bars + 4 * EN_NS, *rempty = bars + 4 * EN_NS + 16;
  constexpr int NRAW = U8 ? L::NRAW8 : EN_NS;              // raw-count ring depth
  constexpr uint32_t RAWSZ = U8 ? TC_NT * EN_RAW8_ROW : EN_RAW1;
  static_assert(L::NRAW8 * TC_NT * EN_RAW8_ROW <= EN_NS * EN_RAW1 && L::NRAW8 <= 16, "uint8 raw ring");
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(smem + L::TMEM_PTR);
  uint32_t* s_lut = reinterpret_cast<uint32_t*>(smem + L::LUT);
  const float* x = reinterpret_cast<const float*>(xv);
  const uint8_t* x8 = reinterpret_cast<const uint8_t*>(xv);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int64_t g0 = (int64_t)blockIdx.x * TC_M;
  const int nR = (int)((B + TC_NT - 1) / TC_NT);
  const int rt0 = blockIdx.y * tiles_per_split;
  const int ntiles = max(0, min(nR, rt0 + tiles_per_split) - rt0);

  if (warp == 1 && lane == 0) {
    // uint8: a 32-row tile is converted by ONE group of 4 warps (group g takes the tiles j % 4 == g, i.e. stage g):
    // four tiles are in conversion at once and a thread pays one proxy fence + one barrier hand-off per 32 counts
    for (int i = 0; i < EN_NS; ++i) {
      tc::mbar_init(&dfull[i], 1);
      tc::mbar_init(&xfull[i], U8 ? 4 : EN_CONV_WARPS);
      tc::mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < NRAW; ++i) {
      tc::mbar_init(&rfull[i], 32 * EN_LOAD_WARPS);
      tc::mbar_init(&rempty[i], U8 ? 4 : EN_CONV_WARPS);
    }
    tc::mbar_init(accfull, 1);
    tc::fence_mbar_init();
  }
  if (warp == 2) tc::tmem_alloc(s_tmem, 128);
  if (U8) en_build_lut(s_lut, log1p_in, tid, EN_THREADS);
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem_base = *s_tmem;
  const int lid = en_loader_id(warp);

  if (lid >= 0 && U8) {
    // ===== raw producers, uint8 counts: a row's 128 genes are 8 x 16 B; one warp instruction copies 4 rows
    const uint32_t rowbytes = (uint32_t)min((int64_t)TC_M, G - g0);
    const int piece = lane & 7, rsub = lane >> 3;
    const bool lane_on = (uint32_t)piece * 16 < rowbytes;
    auto row_id = [&](int j) -> int64_t {
      const int64_t b = (int64_t)(rt0 + j) * TC_NT + lane;
      return (j < ntiles && b < B) ? (row_idx ? row_idx[b] : b) : -1;
    };
    int64_t rid = row_id(0);
    for (int j = 0; j < ntiles; ++j) {
      const int s = j % NRAW;
      const int64_t ridn = row_id(j + 1);
      tc::mbar_wait(&rempty[s], ((j / NRAW) & 1) ^ 1);
      const uint32_t dst0 = tc::smem_u32(smem + L::RAW + s * RAWSZ) + piece * 16;
#pragma unroll
      for (int i = 0; i < 8 / EN_LOAD_WARPS; ++i) {
        const int rr = 4 * (lid + EN_LOAD_WARPS * i) + rsub;
        const int64_t r = __shfl_sync(0xffffffffu, rid, rr);
        if (r >= 0 && lane_on) tc::cp_async16(dst0 + rr * EN_RAW8_ROW, x8 + r * ldx + g0 + piece * 16);
      }
      tc::cp_async_mbar_arrive_noinc(&rfull[s]);
      rid = ridn;
    }
  } else if (lid >= 0) {
    // ===== raw producers: per 32-row tile, loader warp `lid` copies the 512-byte gene segments of rows
    // lid, lid + 4, ... (one warp instruction = one row segment); lane l holds the id of the tile's row l
    const uint32_t rowbytes = (uint32_t)min((int64_t)TC_M, G - g0) * 4;
    const bool lane_on = (uint32_t)lane * 16 < rowbytes;
    auto row_id = [&](int j) -> int64_t {
      const int64_t b = (int64_t)(rt0 + j) * TC_NT + lane;
      return (j < ntiles && b < B) ? (row_idx ? row_idx[b] : b) : -1;
    };
    int64_t rid = row_id(0);
    for (int j = 0; j < ntiles; ++j) {
      const int s = j % EN_NS;
      const int64_t ridn = row_id(j + 1);
      tc::mbar_wait(&rempty[s], ((j / EN_NS) & 1) ^ 1);
      const uint32_t dst0 = tc::smem_u32(smem + L::RAW + s * EN_RAW1) + lane * 16;
#pragma unroll
      for (int i = 0; i < TC_NT / EN_LOAD_WARPS; ++i) {
        const int rr = lid + EN_LOAD_WARPS * i;
        const int64_t r = __shfl_sync(0xffffffffu, rid, rr);
        if (r >= 0 && lane_on) tc::cp_async16(dst0 + rr * EN_RAW2_ROW, x + r * ldx + g0 + lane * 4);
      }
      tc::cp_async_mbar_arrive_noinc(&rfull[s]);
      rid = ridn;
    }
  } else if (warp == 3) {
    // ===== dy producer: one packed 32-row tile per stage (bulk copy)
    if (lane == 0) {
      for (int j = 0; j < ntiles; ++j) {
        const int s = j % EN_NS;
        tc::mbar_wait(&empty[s], ((j / EN_NS) & 1) ^ 1);
        tc::mbar_arrive_expect_tx(&dfull[s], 2 * TC_BLK32);
        tc::bulk_g2s(smem + L::A + s * 2 * EN_BLK, dypack + (size_t)(rt0 + j) * 2 * TC_BLK32, 2 * TC_BLK32, &dfull[s]);
      }
    }
  } else if (warp == 1) {
    if (ntiles > 0) {  // warp-uniform issue loop, one elected lane issues (tc.cuh)
      // D[128 hidden, 128 genes] += dy^T (MN-major A) * log1p(x)^T (MN-major B); K = 32 rows per tile
      constexpr uint32_t idesc = tc::idesc_f16_f32(TC_M, TC_M, 1, 1);
      const uint32_t leader = tc::elect_one();
      // stage 0 operands (stage s = + s * 2 * EN_BLK): MN-major 32 x 128 blocks: LBO = 128, SBO = 32*16, K-step 256 B
      const uint32_t a0 = tc::smem_u32(smem + L::A), b0 = tc::smem_u32(smem + L::W);
      const tc::Oper A0 = tc::make_oper(a0, a0 + TC_BLK32, 128, TC_NT * 16, 256);
      const tc::Oper B0 = tc::make_oper(b0, b0 + TC_BLK32, 128, TC_NT * 16, 256);
      for (int j = 0; j < ntiles; ++j) {
        const int s = j % EN_NS;
        tc::mbar_wait(&dfull[s], (j / EN_NS) & 1);
        tc::mbar_wait(&xfull[s], (j / EN_NS) & 1);
        tc::fence_after_sync();
        if (leader) {
          tc::Oper a = A0, b = B0;
          a.hi_lo += s * ((2 * EN_BLK) >> 4); a.lo_lo += s * ((2 * EN_BLK) >> 4);
          b.hi_lo += s * ((2 * EN_BLK) >> 4); b.lo_lo += s * ((2 * EN_BLK) >> 4);
          tc::mma3<TC_NT / 16>(tmem_base, a, b, idesc, j > 0);
          tc::mma_commit(&empty[s]);
        }
        __syncwarp();
      }
      if (leader) tc::mma_commit(accfull);
      __syncwarp();
    }
  } else if (warp >= 4) {
    // ===== converters: thread = (row r of the 32-row tile = lane, chunk gc of 8 genes = warp - 4)
    const int gc = warp - 4;
    const int64_t gb = g0 + 8 * gc;
    const float sX = tc::pow2i(EN_X_SCALE_LOG2);
    static_assert(EN_NS == 4 && EN_CONV_WARPS == 16, "uint8 converter groups");
    for (int j = (warp - 4) >> 2; U8 && j < ntiles; j += 4) {
      const int s = j % EN_NS, rs = j % NRAW, wq = (warp - 4) & 3;  // this warp: genes 32 wq .. 32 wq + 31 of row `lane`
      const bool rvalid = (int64_t)(rt0 + j) * TC_NT + lane < B;
      tc::mbar_wait(&rfull[rs], (j / NRAW) & 1);
      const uint8_t* raw = smem + L::RAW + rs * RAWSZ + lane * EN_RAW8_ROW + wq * 32;
      uint4 b0 = *reinterpret_cast<const uint4*>(raw);
      uint4 b1 = *reinterpret_cast<const uint4*>(raw + 16);
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(&rempty[rs]);
      // rows past the end of the batch were not copied (they are K entries of the product: must be zero); genes
      // past G only reach output columns that are never stored
      if (!rvalid) b0 = b1 = make_uint4(0u, 0u, 0u, 0u);
      uint4 hi[4], lo[4];
      en_lut8(s_lut, b0.x, b0.y, hi[0], lo[0]);
      en_lut8(s_lut, b0.z, b0.w, hi[1], lo[1]);
      en_lut8(s_lut, b1.x, b1.y, hi[2], lo[2]);
      en_lut8(s_lut, b1.z, b1.w, hi[3], lo[3]);
      tc::mbar_wait(&empty[s], ((j / EN_NS) & 1) ^ 1);
      uint8_t* o = smem + L::W + s * 2 * EN_BLK + (4 * wq) * (TC_NT * 16) + lane * 16;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        *reinterpret_cast<uint4*>(o + k * (TC_NT * 16)) = hi[k];
        *reinterpret_cast<uint4*>(o + k * (TC_NT * 16) + TC_BLK32) = lo[k];
      }
      tc::fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(&xfull[s]);
    }
    for (int j = 0; !U8 && j < ntiles; ++j) {
      const int s = j % EN_NS;
      const bool rvalid = (int64_t)(rt0 + j) * TC_NT + lane < B;
      tc::mbar_wait(&rfull[s], (j / EN_NS) & 1);
      const uint8_t* raw = smem + L::RAW + s * EN_RAW1 + lane * EN_RAW2_ROW + gc * 32;
      const float4 x0 = *reinterpret_cast<const float4*>(raw);
      const float4 x1 = *reinterpret_cast<const float4*>(raw + 16);
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(&rempty[s]);
      float v[8] = {x0.x, x0.y, x0.z, x0.w, x1.x, x1.y, x1.z, x1.w};
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float xv = (rvalid && gb + i < G) ? v[i] : 0.f;
        v[i] = (log1p_in ? fm::flog(1.f + xv) : xv) * sX;
        v[i] = fminf(fmaxf(v[i], -60000.f), 60000.f);
      }
      uint4 hi, lo;
      tc::split8(v, hi, lo);
      tc::mbar_wait(&empty[s], ((j / EN_NS) & 1) ^ 1);
      uint8_t* o = smem + L::W + s * 2 * EN_BLK + gc * (TC_NT * 16) + lane * 16;
      *reinterpret_cast<uint4*>(o) = hi;
      *reinterpret_cast<uint4*>(o + TC_BLK32) = lo;
      tc::fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(&xfull[s]);
    }
    if (ntiles > 0) {
      tc::mbar_wait(accfull, 0);
      tc::fence_after_sync();
    }
    // epilogue: thread = (hidden n = 32 * (warp % 4) + lane, gene group cg of 32 genes)
    const int q = warp & 3, cg = (warp - 4) >> 2;
    const int n = 32 * q + lane;
    const uint32_t lane_sel = (uint32_t)(32 * q) << 16;
    const int e = max(-100, min(100, (int)ceilf(log2f(fmaxf(*absmax, 1e-30f)))));
    const float k = tc::pow2i(e - 14) * tc::pow2i(-EN_X_SCALE_LOG2);
    float* o = slab + ((int64_t)blockIdx.y * TC_M + n) * G + g0 + 32 * cg;
#pragma unroll 1
    for (int c = 0; c < 4; ++c) {
      float v[8];
      if (ntiles > 0) {
        tc::tmem_ld8(tmem_base + lane_sel + 32 * cg + 8 * c, v);
        tc::tmem_ld_wait();
      } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] = 0.f;
      }
      // 8 consecutive genes of one slab row: two 16-byte stores when the row is aligned and the chunk inside G
      if (xvec && g0 + 32 * cg + 8 * c + 8 <= G) {
        *reinterpret_cast<float4*>(o + 8 * c) = make_float4(v[0] * k, v[1] * k, v[2] * k, v[3] * k);
        *reinterpret_cast<float4*>(o + 8 * c + 4) = make_float4(v[4] * k, v[5] * k, v[6] * k, v[7] * k);
      } else {
#pragma unroll
        for (int i = 0; i < 8; ++i)
          if (g0 + 32 * cg + 8 * c + i < G) o[8 * c + i] = v[i] * k;
      }
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 2) tc::tmem_dealloc(tmem_base, 128);
}

// ---------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------
struct EnFwdPlan {
  int64_t nKT, rowtiles;          // nKT = packed 32-gene K-tiles (whole 128-gene chunks)
  int nsplit, tiles_per_split;    // split-K in 128-gene chunks
  size_t off_w, off_y, off_rs, total;
};
static EnFwdPlan en_fwd_plan(int64_t B, int64_t K) {
  EnFwdPlan p;
  const int64_t nKC = cdiv(K, TC_M);
  p.nKT = 4 * nKC;
  p.rowtiles = cdiv(B, TC_M);
  int64_t ns = std::min<int64_t>(std::max<int64_t>(1, 148 / std::max<int64_t>(1, p.rowtiles)), std::max<int64_t>(1, nKC / 2));
  p.tiles_per_split = (int)cdiv(nKC, ns);
  p.nsplit = (int)cdiv(nKC, p.tiles_per_split);
  auto al = [](size_t v) { return (v + 255) / 256 * 256; };
  p.off_w = 0;
  p.off_y = al((size_t)p.nKT * 2 * EN_BLK);
  p.off_rs = al(p.off_y + (size_t)p.nsplit * B * TC_H * sizeof(float));
  p.total = al(p.off_rs + (size_t)p.nsplit * B * sizeof(float));
  return p;
}
int64_t tc_enc_fwd_ws_bytes(int64_t B, int64_t K) { return (int64_t)en_fwd_plan(B, K).total; }

int tc_enc_fwd(const void* x, int64_t ldx, const int64_t* row_idx, int64_t B, int64_t K, const float* W, int64_t ldw,
               const float* bias, int flags, float* y, int64_t ldy, float* library, float* bn_partials, void* ws,
               cudaStream_t st) {
  const bool u8 = (flags & SCVI_LIN_X_U8) != 0;
  SCVI_REQUIRE(ws != nullptr && ((uintptr_t)ws % 256 == 0), "scvi_linear_fwd(SCVI_LIN_TC): ws must be 256-byte aligned");
  const EnFwdPlan p = en_fwd_plan(B, K);
  uint8_t* w8 = (uint8_t*)ws;
  uint8_t* wpack = w8 + p.off_w;
  float* ypart = (float*)(w8 + p.off_y);
  float* rs = (float*)(w8 + p.off_rs);
  tc_pack_we_kernel<<<(unsigned)cdiv(p.nKT * 4 * TC_M, 256), 256, 0, st>>>(W, ldw, K, p.nKT, wpack);
  SCVI_LAUNCH_CHECK("scvi_linear_fwd(tc pack)");
  static std::atomic<uint64_t> optin[2];
  auto kern = u8 ? tc_enc_fwd_kernel<true> : tc_enc_fwd_kernel<false>;
  const int smem_bytes = (int)(u8 ? En1SmemU8::TOTAL : En1Smem::TOTAL) + 128;
  if (int rc = smem_optin((const void*)kern, smem_bytes, optin[u8], "scvi_linear_fwd")) return rc;
  dim3 grid((unsigned)p.rowtiles, (unsigned)p.nsplit);
  kern<<<grid, EN_THREADS, smem_bytes, st>>>(x, ldx, row_idx, B, K, wpack, p.tiles_per_split,
                                                     (flags & SCVI_LIN_LOG1P) ? 1 : 0, ypart, rs);
  SCVI_LAUNCH_CHECK("scvi_linear_fwd(tc)");
  tc_enc_fwd_combine_kernel<<<(unsigned)cdiv(B, SCVI_BN_TILE), 512, 0, st>>>(
      ypart, rs, p.nsplit, B, bias, y, ldy, (flags & SCVI_LIN_LIBRARY) ? library : nullptr, bn_partials);
  SCVI_LAUNCH_CHECK("scvi_linear_fwd(tc combine)");
  return SCVI_OK;
}

constexpr int EN_ABSMAX_BLOCKS = 148 * 8;
struct EnDwPlan {
  int64_t nT, nR;
  int nsplit, tiles_per_split;
  size_t off_dy, off_scal, off_slab, total;
};
static EnDwPlan en_dw_plan(int64_t B, int64_t K) {
  EnDwPlan p;
  p.nT = cdiv(K, TC_M);
  p.nR = cdiv(B, TC_NT);
  int64_t ns = std::min<int64_t>(std::max<int64_t>(1, 148 / p.nT), std::max<int64_t>(1, p.nR / 4));
  p.tiles_per_split = (int)cdiv(p.nR, ns);
  p.nsplit = (int)cdiv(p.nR, p.tiles_per_split);
  auto al = [](size_t v) { return (v + 255) / 256 * 256; };
  p.off_dy = 0;
  p.off_scal = al((size_t)p.nR * 2 * TC_BLK32);
  p.off_slab = al(p.off_scal + (EN_ABSMAX_BLOCKS + 8) * sizeof(float));
  p.total = al(p.off_slab + (size_t)p.nsplit * TC_M * K * sizeof(float));
  return p;
}
int64_t tc_enc_dw_ws_bytes(int64_t B, int64_t K) { return (int64_t)en_dw_plan(B, K).total; }

// dW[128, K] (row stride ldw == K) = dy^T f(x)
int tc_enc_dw(const void* x, int64_t ldx, const int64_t* row_idx, int64_t B, int64_t K, const float* dy, int64_t lddy,
              int flags, float* dW, const float* dy_absmax_part, int64_t n_absmax_part, void* ws, cudaStream_t st) {
  const bool u8 = (flags & SCVI_LIN_X_U8) != 0;
  SCVI_REQUIRE(ws != nullptr && ((uintptr_t)ws % 256 == 0), "scvi_linear_bwd(SCVI_LIN_TC): ws must be 256-byte aligned");
  SCVI_REQUIRE(!dy_absmax_part || n_absmax_part > 0, "scvi_linear_bwd(SCVI_LIN_TC): empty dy_absmax_part");
  SCVI_REQUIRE(lddy % 4 == 0 && ((uintptr_t)dy % 16 == 0), "scvi_linear_bwd(SCVI_LIN_TC): dy must be 16-byte aligned");
  const EnDwPlan p = en_dw_plan(B, K);
  uint8_t* w8 = (uint8_t*)ws;
  uint8_t* dypack = w8 + p.off_dy;
  float* scal = (float*)(w8 + p.off_scal);  // [0] = max |dy|, [8...] = per-block partial maxima
  float* slab = (float*)(w8 + p.off_slab);
  if (!dy_absmax_part) {
    const int nb = (int)std::min<int64_t>(cdiv(B * (TC_H / 4), 256), EN_ABSMAX_BLOCKS);
    tc_absmax_kernel<<<nb, 256, 0, st>>>(dy, lddy, B, TC_H, scal + 8);
    SCVI_LAUNCH_CHECK("scvi_linear_bwd(tc absmax)");
    dy_absmax_part = scal + 8;
    n_absmax_part = nb;
  }
  tc_pack_rows32_kernel<<<(unsigned)p.nR, 512, 0, st>>>(dy, lddy, B, dy_absmax_part, (int)n_absmax_part, scal, dypack);
  SCVI_LAUNCH_CHECK("scvi_linear_bwd(tc pack)");
  static std::atomic<uint64_t> optin[2];
  auto kern = u8 ? tc_enc_dw_kernel<true> : tc_enc_dw_kernel<false>;
  if (int rc = smem_optin((const void*)kern, (int)EnSmem::TOTAL + 128, optin[u8], "scvi_linear_bwd")) return rc;
  const int xvec = (K % 4 == 0) && (u8 || ((ldx % 4 == 0) && ((uintptr_t)x % 16 == 0)));  // also: slab rows 16-byte aligned
  dim3 grid((unsigned)p.nT, (unsigned)p.nsplit);
  kern<<<grid, EN_THREADS, EnSmem::TOTAL + 128, st>>>(x, ldx, row_idx, B, K, dypack, scal, p.tiles_per_split,
                                                    (flags & SCVI_LIN_LOG1P) ? 1 : 0, xvec, slab);
  SCVI_LAUNCH_CHECK("scvi_linear_bwd(tc dW)");
  return launch_colsum(slab, p.nsplit, TC_M * K, TC_M * K, colsum_seg1(dW, TC_M * K), st, "scvi_linear_bwd(tc reduce)");
}

}  // namespace scvi
