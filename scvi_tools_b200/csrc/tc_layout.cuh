// Operand layouts shared by the tcgen05 decoder kernels.
//
// Every fp16 operand block is stored in the UMMA no-swizzle "core matrix" layout
//     addr(r, c) = c * (R * 16 B) + r * 16 B        r < R rows, c = 16-byte chunk of 8 halves
// which one buffer serves twice:
//   * as a K-major operand (rows = M or N index, chunks run along the reduction):  LBO = R*16, SBO = 128
//   * as an MN-major operand (rows = reduction index, chunks run along M or N):    LBO = 128,  SBO = R*16
// so the weight tile that multiplies h in the logits GEMM is also the B operand of dh = ds * W, and the h
// tile that multiplies W is also the B operand of dW = ds^T * h -- no transposed copies.
//
// fp32 accuracy comes from an error-compensated fp16 split v = hi + lo (2 x 11-bit significands):
// D = hi*hi + hi*lo + lo*hi (3 MMAs per K-step).  Operands are pre-scaled by powers of two so the `lo`
// parts stay in fp16's normal range; accumulators are rescaled in the epilogue.
#pragma once
#include "fastmath.cuh"
#include "tc.cuh"

namespace scvi {

constexpr int TC_M = 128;        // UMMA M
constexpr int TC_N = 128;        // genes per tile of the LSE / NLL-only kernels
constexpr int TC_H = 128;        // n_hidden (K of the logits GEMM)
constexpr int TC_KC = TC_H / 8;  // 16-byte chunks per row
constexpr float TC_SCALE_H = 4.f, TC_SCALE_W = 256.f, TC_RESCALE = 1.f / (4.f * 256.f);
constexpr uint32_t TC_OPER_BYTES = TC_M * TC_H * 2;     // one 128 x 128 fp16 block: 32 KB
constexpr uint32_t TC_STAGE_BYTES = 2 * TC_OPER_BYTES;  // hi + lo
constexpr int TC_GC = 8;  // floats per gene in the constant table:
                          // theta, log(theta+eps), lgamma(theta), bs, bp, 1/(theta+eps), digamma(theta), 0

// floats per gene of the gconst workspace region: [constants TC_GC][contiguous scale bias 1][dgd table DGT_LD]
// (three consecutive arrays over the padded gene count, not interleaved)
constexpr int TC_GCONST_FLOATS = TC_GC + 1 + fm::DGT_LD;

// 32-row blocks (narrow tiles of the training kernels): one fp16 block 32 x 128 = 8 KB
constexpr int TC_NT = 32;
constexpr uint32_t TC_BLK32 = TC_NT * TC_H * 2;
constexpr uint32_t TC_GC32_BYTES = TC_NT * TC_GC * 4;  // gene constants of one 32-gene tile: 1 KB
constexpr uint32_t TC_LGT32_BYTES = TC_NT * fm::LGT * 4;  // lgamma-difference tables of one 32-gene tile: 8 KB
// packed32 tile (global): [Ws hi][Ws lo]([Wp hi][Wp lo])[gene constants][lgd tables]
__host__ __device__ constexpr uint32_t tc_w32_stage_bytes(int nheads) {
  return nheads * 2 * TC_BLK32 + TC_GC32_BYTES + TC_LGT32_BYTES;
}

constexpr int TC_ROWC = 8;  // floats per row in the row-constant table of the dW kernel
constexpr uint32_t TC_H32_STAGE_BYTES = 2 * TC_BLK32 + TC_NT * TC_ROWC * 4;  // [h hi][h lo][row constants]: 17 KB

// staged-operand scales of the training kernels (powers of two)
constexpr int TC_P_SCALE_LOG2 = 14;   // softmax probabilities p in [0,1]
constexpr int TC_DP_SCALE_LOG2 = 10;  // dropout-logit gradients in [-1,1]

__device__ __forceinline__ fm::GeneK load_gene_k(const float* __restrict__ gc) {
  fm::GeneK g;
  g.theta = gc[0];
  g.log_theta = gc[1];
  g.lg_theta = gc[2];
  g.rth = gc[5];
  g.dg_theta = gc[6];
  return g;
}

}  // namespace scvi
