/*
 * scvi_b200.h -- C ABI of the B200-native SCVI VAE minibatch training hot path.
 *
 * The reference (scverse/scvi-tools v1.5.0.post1) is pure Python/PyTorch and has NO FFI for
 * this path; its boundary is the Python class surface.  This header is
 * the ABI a maintainer binds UNDER that surface (ctypes stub in INTEGRATION.md): each entry
 * point names the reference code it replaces (paths relative to the reference root).
 *
 * Conventions
 *  - every pointer is a DEVICE pointer unless the name ends in `_host`; row-major; fp32
 *    values, int64 row / category indices (the reference's dtypes: src/scvi/data/_utils.py:37-47)
 *  - `stream` is a `cudaStream_t` passed as `void*` (NULL = legacy default stream)
 *  - every function returns 0 on success, a negative SCVI_E* code otherwise; it never
 *    throws, never exits, never synchronises the device; `scvi_last_error_string()` holds the
 *    message of the calling thread's last failure
 *  - the library allocates nothing persistent: all buffers (workspaces too) are caller-owned
 *  - "BN tile" = SCVI_BN_TILE consecutive rows; kernels that produce a pre-BatchNorm
 *    activation also emit per-tile column statistics (mean, M2) merged later by
 *    scvi_bn_finalize (Chan's parallel variance; deterministic)
 */
#ifndef SCVI_B200_H
#define SCVI_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SCVI_OK 0
#define SCVI_EINVAL (-1)   /* bad argument (shape, alignment, NULL) */
#define SCVI_ECUDA (-2)    /* CUDA runtime error at launch */
#define SCVI_EUNSUPPORTED (-3) /* configuration outside what the kernels implement */

#define SCVI_BN_TILE 64

/* likelihood selector: reference `gene_likelihood` (src/scvi/module/_vae.py:520-528) */
#define SCVI_LIK_NB 0
#define SCVI_LIK_ZINB 1

/* scvi_linear_fwd / _bwd flags */
#define SCVI_LIN_LOG1P 1   /* input is log1p(x)   (src/scvi/module/_vae.py:373-374) */
#define SCVI_LIN_LIBRARY 2 /* also emit library = log(sum_g x)  (src/scvi/module/_vae.py:371-372) */
#define SCVI_LIN_TC 4      /* tcgen05 tensor-core path (error-compensated fp16 split, fp32-grade): needs
                              N == 128, SCVI_LIN_LOG1P, no covariates, and a workspace */

/* GEMM arithmetic of the fused decoder kernels */
#define SCVI_PREC_FP32 0   /* FFMA, fp32 everywhere: the parity path */
#define SCVI_PREC_TC 1     /* tcgen05 tensor-core path (when built) */
#define SCVI_PREC_PREPACKED 0x100 /* OR-ed into `precision` of scvi_decoder_nll_fwd: the workspace already holds the
                                     packed weight tiles (scvi_decoder_nll_prepack ran on the current weights) */
/* Count operand of the tensor-core training kernels (OR-ed into `precision` of scvi_decoder_nll_fwd / _bwd and
 * into `flags` of scvi_linear_fwd / _bwd as SCVI_LIN_X_INTEGER / SCVI_LIN_X_U8):
 *   SCVI_X_INTEGER  the fp32 count matrix holds non-negative integers only (what AnnData `X` / a counts layer
 *                   holds; the data classes verify it once at load time) -- enables the straight-line per-element
 *                   path (table look-ups, no float<->int conversions);
 *   SCVI_X_U8       `x` points to uint8 counts (all counts < 256; ldx in bytes, rows 8-byte aligned for vector
 *                   loads): 1 byte per count from HBM.  Training kernels only (SCVI_PREC_TC with dh / dW). */
#define SCVI_X_INTEGER 0x200
#define SCVI_X_U8 0x400
/* SCVI_PREC_AOPS (with SCVI_PREC_TC and an integer count kind, OR-ed into `precision` of scvi_decoder_nll_fwd_ws_bytes,
 * scvi_decoder_nll_fwd with dh, and the matching scvi_decoder_nll_bwd with fwd_ws): the fused likelihood kernel keeps
 * its staged MMA operands (fp16 hi/lo, in the forward workspace) and the weight-gradient kernel becomes a tensor-core
 * GEMM over them instead of recomputing the logits and every element (measured slower than the recompute kernel on
 * B200: opt-in, see experiments/README.md). */
#define SCVI_PREC_AOPS 0x800
#define SCVI_LIN_X_U8 SCVI_X_U8 /* scvi_linear_fwd / _bwd with SCVI_LIN_TC: `x` points to uint8 counts, ldx in bytes,
                                   rows 16-byte aligned; the converters look log1p(count) up in a 256-entry table
                                   built from the fp32 path's own arithmetic (results bit-identical to fp32 counts) */

int scvi_b200_abi_version(void); /* 3 */
const char* scvi_last_error_string(void);
/* number of kernels this library has launched (or recorded into a stream capture) since load */
int64_t scvi_launch_count(void);
/* Per-kernel event timing of eager launches on `stream` (bench.py's roofline leg; never inside a stream
 * capture).  _begin queues a `delay_us` spin kernel so the host runs ahead and the kernels execute
 * back to back; _end synchronises and writes "label\tlaunches\ttotal_ms\n" lines (labels = the names in
 * this library's launch checks), returning the text length or -1. */
int scvi_profile_begin(void* stream, int64_t delay_us);
int64_t scvi_profile_end(char* buf, int64_t cap);

/* ---- a1: minibatch row gather --------------------------------------------------------
 * replaces AnnTorchDataset.__getitem__ (src/scvi/data/_anntorchdataset.py:115-169,
 * `data[indexes].astype(float32)`) + the Lightning host->device copy, and
 * _DeviceBackedDataset.__getitem__ (src/scvi/dataloaders/_data_splitting.py:589-732).
 * out[b, :] = X[idx[b], :] bit-exactly.  idx values must lie in [0, n_rows). */
int scvi_gather_rows(const float* X, int64_t n_rows, int64_t G, int64_t ldx, const int64_t* idx,
                     int64_t B, float* out, int64_t ldo, void* stream);
/* out[b] = src[idx[b]] for the int64 `batch` / `labels` columns (shape (B,1) in the reference) */
int scvi_gather_i64(const int64_t* src, int64_t n_rows, const int64_t* idx, int64_t B, int64_t* out,
                    void* stream);
/* same gather from a compact integer matrix -- resident in HBM, or a minibatch staged from pinned host
 * memory -- in the narrowest lossless count type (uint16: counts < 65536; uint8: counts < 256), densified
 * to fp32 exactly (`data[indexes].astype(float32)`, src/scvi/data/_anntorchdataset.py:143-147, for an
 * integer-typed `adata.X`).  idx == NULL means the identity (expand rows 0..B-1 in place order). */
int scvi_gather_rows_u16(const uint16_t* X, int64_t n_rows, int64_t G, int64_t ldx,
                         const int64_t* idx, int64_t B, float* out, int64_t ldo, void* stream);
/* HOST code (no CUDA call): out[b, :] = X[idx[b], :] for a matrix in host memory, rows of `row_bytes` bytes, on a
 * persistent pool of `n_threads` host threads (0 = all hardware threads) -- the per-minibatch `data[indexes]` of
 * src/scvi/data/_anntorchdataset.py:143-152 ahead of the H2D copy.  Blocking; idx is checked against n_rows. */
int scvi_host_gather_rows(const void* X, int64_t n_rows, int64_t row_bytes, int64_t ldx_bytes, const int64_t* idx,
                          int64_t B, void* out, int64_t ldo_bytes, int n_threads);
/* uint8 rows copied unexpanded (for the kernels that read uint8 counts, SCVI_X_U8): out[b, :] = X[idx[b], :] */
int scvi_gather_rows_u8_raw(const uint8_t* X, int64_t n_rows, int64_t G, int64_t ldx, const int64_t* idx, int64_t B,
                            uint8_t* out, int64_t ldo, void* stream);
int scvi_gather_rows_u8(const uint8_t* X, int64_t n_rows, int64_t G, int64_t ldx,
                        const int64_t* idx, int64_t B, float* out, int64_t ldo, void* stream);

/* same gather from a CSR matrix resident in HBM (sparse `adata.X`: `data[indexes].toarray()`,
 * src/scvi/data/_anntorchdataset.py:143-152; torch-sparse conversion src/scvi/data/_utils.py:50-87):
 * indptr int64 [n_rows+1]; indices uint16 (index_bytes 2, G <= 65536) or int32 (4); values float32
 * (value_kind 0), uint16 (1) or uint8 (2).  Rows are densified to fp32 exactly; duplicates within a
 * row are not summed (canonical CSR expected).  Column ids outside [0, G) are skipped and counted in
 * *n_bad (device counter, nullable). */
int scvi_gather_rows_csr(const int64_t* indptr, const void* indices, int index_bytes, const void* values,
                         int value_kind, int64_t n_rows, int64_t G, const int64_t* idx, int64_t B,
                         float* out, int64_t ldo, unsigned long long* n_bad, void* stream);

/* ---- a7/a8: standalone likelihood kernels ("NB-loss kernel") --------------------------
 * replace log_nb_positive / log_zinb_positive
 * (src/scvi/distributions/_negative_binomial.py:99-135, :36-96) as called by
 * NegativeBinomial.log_prob (:445-459) / ZeroInflatedNegativeBinomial.log_prob (:573-586).
 * x, mu, pi, out are [B,G] with row strides ld*; theta is [G] (theta_ld = 0) or [B,G].
 * eps is the reference's 1e-8.  `bad_support` (nullable, int32[1]) is set to 1 if any x is
 * negative or non-integer -- the lazily raised counterpart of _validate_sample (:575-582). */
int scvi_nb_logprob_fwd(const float* x, int64_t ldx, const float* mu, int64_t ldmu, const float* theta,
                        int64_t theta_ld, int64_t B, int64_t G, float eps, float* out, int64_t ldo,
                        int32_t* bad_support, void* stream);
int scvi_zinb_logprob_fwd(const float* x, int64_t ldx, const float* mu, int64_t ldmu,
                          const float* theta, int64_t theta_ld, const float* pi, int64_t ldpi,
                          int64_t B, int64_t G, float eps, float* out, int64_t ldo,
                          int32_t* bad_support, void* stream);
/* backward: gout is d(loss)/d(out) [B,G]; writes dmu, (dpi) [B,G] and dtheta either [B,G]
 * (theta_ld != 0) or per-row-tile partials [ceil(B/SCVI_BN_TILE), G] to be summed by
 * scvi_colsum_partials (theta_ld == 0).  Any of dmu/dtheta/dpi may be NULL (skipped). */
int scvi_nb_logprob_bwd(const float* x, int64_t ldx, const float* mu, int64_t ldmu, const float* theta,
                        int64_t theta_ld, const float* gout, int64_t ldg, int64_t B, int64_t G,
                        float eps, float* dmu, float* dtheta, void* stream);
int scvi_zinb_logprob_bwd(const float* x, int64_t ldx, const float* mu, int64_t ldmu,
                          const float* theta, int64_t theta_ld, const float* pi, int64_t ldpi,
                          const float* gout, int64_t ldg, int64_t B, int64_t G, float eps, float* dmu,
                          float* dtheta, float* dpi, void* stream);
/* ---- negative-binomial mixture likelihood (TOTALVI proteins) ----------
 * replaces log_mixture_nb with a shared inverse dispersion (theta_2 = None) as called by
 * NegativeBinomialMixture.log_prob (src/scvi/distributions/_negative_binomial.py:138-216, :589-717):
 *   out = logsumexp(log NB(x; mu1, theta), log NB(x; mu2, theta) - pi) - softplus(-pi),  pi = mixture logit.
 * x, mu1, mu2, pi, out are [B,P] with row strides ld*; theta is [P] (theta_ld = 0) or [B,P].
 * backward: same conventions as scvi_nb_logprob_bwd (dtheta per-row-tile partials when theta_ld == 0). */
int scvi_mixture_nb_logprob_fwd(const float* x, int64_t ldx, const float* mu1, int64_t ldmu1, const float* mu2,
                                int64_t ldmu2, const float* theta, int64_t theta_ld, const float* pi,
                                int64_t ldpi, int64_t B, int64_t P, float eps, float* out, int64_t ldo,
                                int32_t* bad_support, void* stream);
int scvi_mixture_nb_logprob_bwd(const float* x, int64_t ldx, const float* mu1, int64_t ldmu1, const float* mu2,
                                int64_t ldmu2, const float* theta, int64_t theta_ld, const float* pi,
                                int64_t ldpi, const float* gout, int64_t ldg, int64_t B, int64_t P, float eps,
                                float* dmu1, float* dmu2, float* dtheta, float* dpi, void* stream);
/* out[n] = sum_t partials[t, n] in fixed order (deterministic) */
int scvi_colsum_partials(const float* partials, int64_t n_tiles, int64_t N, float* out, int accumulate,
                         void* stream);

/* ---- a2/a3: FCLayers linear layer with covariate injection ----------------------------
 * replaces, per layer, `torch.cat((x, one_hot(cat)...))` + nn.Linear
 * (src/scvi/nn/_base_components.py:148-188,198-215) and for the encoder input the pre-ops
 * `library = log(x.sum(1))`, `x_ = log1p(x)` (src/scvi/module/_vae.py:370-374).
 *   y[b, n] = sum_k f(x[r(b), k]) * W[n, k] + sum_c W[n, cov_cols[b, c]] + bias[n]
 * with r(b) = row_idx ? row_idx[b] : b (reads straight from the resident count matrix),
 * f = log1p when SCVI_LIN_LOG1P.  The one-hot concat becomes a weight-column gather:
 * cov_cols[b, c] is the absolute column of W (>= K) selected by covariate c of cell b.
 * bn_partials (nullable) [ceil(B/SCVI_BN_TILE), 2, N] receives per-tile (mean, M2) of y. */
int64_t scvi_linear_fwd_ws_bytes(int64_t B, int64_t K, int64_t N, int flags); /* 0 unless SCVI_LIN_TC */
int scvi_linear_fwd(const void* x, int64_t ldx, const int64_t* row_idx, int64_t B, int64_t K, int64_t N,
                    const float* W, int64_t ldw, const float* bias, const int32_t* cov_cols,
                    int n_cov, int flags, float* y, int64_t ldy, float* library, float* bn_partials,
                    void* ws, void* stream);
/* backward of the above.  dW [N, ldw] (all K + covariate columns are written: covariate
 * columns receive the per-category column sums of dy -- the one-hot GEMM as a segmented
 * sum), db [N] (nullable), dx [B,K] (nullable; not available with SCVI_LIN_LOG1P).
 * `ws` is a caller-owned workspace of scvi_linear_bwd_ws_bytes() bytes (split-B partial
 * slabs reduced in fixed order: deterministic, no atomics).  n_w_cols = total columns of W.
 * dy_absmax_part (nullable; used by SCVI_LIN_TC): n_absmax_part floats whose maximum is max |dy|
 * (as produced by scvi_mid_bwd) -- saves the pass over dy that fixes the operand scale. */
int64_t scvi_linear_bwd_ws_bytes(int64_t B, int64_t K, int64_t N, int64_t n_w_cols);
int scvi_linear_bwd(const void* x, int64_t ldx, const int64_t* row_idx, int64_t B, int64_t K, int64_t N,
                    const float* W, int64_t ldw, int64_t n_w_cols, const int32_t* cov_cols, int n_cov,
                    int flags, const float* dy, int64_t lddy, float* dW, float* db, float* dx,
                    int64_t lddx, const float* dy_absmax_part, int64_t n_absmax_part, void* ws,
                    void* stream);

/* ---- a3: BatchNorm1d(momentum .01, eps 1e-3) + ReLU + Dropout --------------------------
 * replaces nn.BatchNorm1d / nn.ReLU / nn.Dropout of FCLayers._build_layer
 * (src/scvi/nn/_base_components.py:102-115).
 * scvi_bn_finalize: training != 0: merge tile partials -> batch mean, rstd = 1/sqrt(var_b+eps),
 *   and running <- (1-momentum)*running + momentum*{mean, unbiased var} (nullable buffers);
 *   training == 0: mean = running_mean, rstd = 1/sqrt(running_var+eps). */
int scvi_bn_finalize(const float* bn_partials, int64_t B, int64_t N, int training, float eps, float momentum,
                     float* running_mean, float* running_var, float* mean, float* rstd, void* stream);
/* standalone tile statistics for a [B,N] activation produced elsewhere */
int scvi_bn_partials(const float* y, int64_t ldy, int64_t B, int64_t N, float* bn_partials, void* stream);
/* a = keep * drop_scale * act(gamma * (y - mean) * rstd + beta); mean/rstd/gamma/beta NULL => no
 * norm; relu = 0 => no activation; drop_mask (uint8 keep-mask, nullable) => no dropout. */
int scvi_bn_act_drop_fwd(const float* y, int64_t ldy, int64_t B, int64_t N, const float* mean,
                         const float* rstd, const float* gamma, const float* beta, int relu,
                         const uint8_t* drop_mask, float drop_scale, float* a, int64_t lda, void* stream);
/* full backward (through the batch statistics when training != 0).  `ws` holds
 * scvi_bn_bwd_ws_bytes() bytes.  dgamma/dbeta nullable. */
int64_t scvi_bn_bwd_ws_bytes(int64_t B, int64_t N);
int scvi_bn_act_drop_bwd(const float* da, int64_t ldda, const float* y, int64_t ldy, int64_t B, int64_t N,
                         const float* mean, const float* rstd, const float* gamma, const float* beta,
                         int relu, const uint8_t* drop_mask, float drop_scale, int training, float* dy,
                         int64_t lddy, float* dgamma, float* dbeta, void* ws, void* stream);
/* Allocation-free variant of scvi_bn_act_drop_bwd.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot allocate device memory (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid cudaMalloc).  Same arguments and results as scvi_bn_act_drop_bwd minus
 * `ws`: the per-tile partial sums live in a scratch array built into the library image
 * (SCVI_BN_STATIC_SCRATCH_BYTES = 526,592 bytes = scvi_bn_bwd_ws_bytes(16384, 256); every GPU that loads the
 * library holds it until the process exits).  Shapes with scvi_bn_bwd_ws_bytes(B, N) beyond that are rejected
 * with SCVI_EINVAL and a message naming both sizes.  The call allocates nothing (no cudaMalloc / cudaFree /
 * allocating library routine): it resolves the scratch's device address once per device (cudaGetSymbolAddress,
 * with the calling thread's current device) and launches the same kernels.  ALL calls on a device share the one
 * array, which keeps the previous call's contents -- every word is rewritten before it is read, so nothing
 * needs clearing -- and therefore two calls that overlap on the GPU (different streams, or graph branches that
 * run concurrently) RACE: serialise them on one stream, or use scvi_bn_act_drop_bwd with separate workspaces. */
#define SCVI_BN_STATIC_SCRATCH_BYTES 526592
int64_t scvi_bn_static_scratch_bytes(void);
int scvi_bn_act_drop_bwd_static(const float* da, int64_t ldda, const float* y, int64_t ldy, int64_t B, int64_t N,
                                const float* mean, const float* rstd, const float* gamma, const float* beta,
                                int relu, const uint8_t* drop_mask, float drop_scale, int training, float* dy,
                                int64_t lddy, float* dgamma, float* dbeta, void* stream);

/* ---- a3 / f4: LayerNorm variant of the above and the batch-embedding gradient ---------------------------
 * scvi_ln_act_drop_fwd / _bwd replace nn.LayerNorm(n_out, elementwise_affine=False) -> nn.ReLU -> nn.Dropout of
 * FCLayers._build_layer with use_layer_norm (src/scvi/nn/_base_components.py:102-115): every ROW of y is normalised
 * with its own mean and biased variance (eps inside the square root; torch's default 1e-5).
 *   a = keep * drop_scale * act((y - mean_row) * rstd_row);   stats [B,2] = (mean, rstd) per row (nullable in fwd),
 * consumed by the backward:  dy = rstd * (g - mean_n g - xhat * mean_n(g * xhat)),  g = da through dropout and ReLU. */
int scvi_ln_act_drop_fwd(const float* y, int64_t ldy, int64_t B, int64_t N, float eps, int relu,
                         const uint8_t* drop_mask, float drop_scale, float* a, int64_t lda, float* stats,
                         void* stream);
int scvi_ln_act_drop_bwd(const float* da, int64_t ldda, const float* y, int64_t ldy, int64_t B, int64_t N,
                         const float* stats, int relu, const uint8_t* drop_mask, float drop_scale, float* dy,
                         int64_t lddy, void* stream);
/* Gradient of `scvi.nn.Embedding` (src/scvi/nn/_embedding.py:29-98, used for batch_representation="embedding",
 * src/scvi/module/_vae.py:218-236,385-396,489-498): dW[e, :] = sum over the cells b with idx[b] == e of g[b, :],
 * rows visited in order (deterministic).  The forward is scvi_gather_rows on the table. */
int scvi_embedding_bwd(const float* g, int64_t ldg, const int64_t* idx, int64_t B, int64_t D, int64_t n_emb,
                       float* dW, void* stream);

/* ---- a4/a9: latent heads, reparameterisation, Normal KL -------------------------------
 * replaces Encoder.forward's heads (src/scvi/nn/_base_components.py:311-317):
 *   q_m = a Wm^T + bm;  q_v = exp(a Wv^T + bv) + var_eps;  z = q_m + sqrt(q_v) * noise
 * and kl_divergence(Normal(q_m, sqrt(q_v)), Normal(0,1)).sum(-1) (src/scvi/module/_vae.py:563-565).
 * L <= 32. */
int scvi_latent_fwd(const float* a, int64_t lda, int64_t B, int64_t H, int64_t L, const float* Wm,
                    const float* bm, const float* Wv, const float* bv, const float* noise, float var_eps,
                    float* q_m, float* q_v, float* z, float* kl, void* stream);
/* upstream grads g_* nullable ([B,L] for q_m/q_v/z, [B] for kl).  Writes da [B,H], dWm/dWv [L,H],
 * dbm/dbv [L].  ws: scvi_latent_bwd_ws_bytes(). */
int64_t scvi_latent_bwd_ws_bytes(int64_t B, int64_t H, int64_t L);
int scvi_latent_bwd(const float* a, int64_t lda, int64_t B, int64_t H, int64_t L, const float* Wm,
                    const float* Wv, const float* noise, float var_eps, const float* q_m,
                    const float* q_v, const float* g_qm, const float* g_qv, const float* g_z,
                    const float* g_kl, float* da, int64_t ldda, float* dWm, float* dbm, float* dWv,
                    float* dbv, void* ws, void* stream);

/* ---- a3/a4/a9 fused: the whole [B, n_hidden] middle of the step in a few launches -------------
 * For the fused training plan (n_hidden == 128, n_latent <= 16, n_latent + n_batch <= 80) these two
 * calls replace the sequence scvi_bn_finalize, scvi_bn_act_drop_fwd, scvi_latent_fwd, scvi_linear_fwd
 * (decoder FCLayers layer 0), scvi_bn_finalize, scvi_bn_act_drop_fwd and its mirror in backward; same
 * math, same reference lines, per-32-row-tile partials merged per cluster of 8 CTAs, then in fixed order by every consumer CTA.
 * y1 / bn_partials1: output of scvi_linear_fwd for the encoder layer.  cov_cols: int32 [B], absolute
 * column of Wd ([H, n_w_cols]) selected by the batch covariate (NULL = none).
 * ws: scvi_mid_ws_bytes() bytes, the SAME buffer for the forward and the backward of one step. */
int64_t scvi_mid_ws_bytes(int64_t B, int64_t L, int64_t n_w_cols);
int scvi_mid_fwd(const float* y1, const float* bn_partials1, int64_t B, int64_t H, int64_t L, const float* gamma1,
                 const float* beta1, float* running_mean1, float* running_var1, const uint8_t* drop_mask,
                 float drop_scale, const float* noise, const float* Wm, const float* bm, const float* Wv,
                 const float* bv, float var_eps, const float* Wd, int64_t n_w_cols, const float* bd,
                 const int32_t* cov_cols, const float* gamma2, const float* beta2, float* running_mean2,
                 float* running_var2, float bn_eps, float momentum, float* a1, float* q_m, float* q_v, float* z,
                 float* kl, float* y2, float* h, float* mean1, float* rstd1, float* mean2, float* rstd2, void* ws,
                 void* stream);
/* dh = d(loss)/d(h); g_kl [B] = d(loss)/d(kl).  Writes dy1 [B,H] (gradient w.r.t. the encoder layer's
 * pre-BatchNorm output) and every parameter gradient of the middle (any may be NULL).
 * dy1_absmax_part (nullable): float[scvi_mid_n_tiles(B)], max |dy1| per row tile -- hand it to
 * scvi_linear_bwd(SCVI_LIN_TC) to skip that call's own pass over dy. */
int64_t scvi_mid_n_tiles(int64_t B);
int scvi_mid_bwd(const float* dh, const float* y2, const float* mean2, const float* rstd2, const float* gamma2,
                 const float* beta2, const float* z, const float* Wd, int64_t n_w_cols, const int32_t* cov_cols,
                 const float* g_kl, const float* q_m, const float* q_v, const float* noise, float var_eps,
                 const float* Wm, const float* Wv, const float* a1, const float* y1, const float* mean1,
                 const float* rstd1, const float* gamma1, const float* beta1, const uint8_t* drop_mask,
                 float drop_scale, int64_t B, int64_t H, int64_t L, float* dy1, float* dgamma2, float* dbeta2,
                 float* dWd, float* dbd, float* dWm, float* dbm, float* dWv, float* dbv, float* dgamma1, float* dbeta1,
                 float* db_enc, float* dy1_absmax_part, void* ws, void* stream);

/* ---- a5/a6/a7/a8/a9: fused decoder heads + softmax + library scaling + NB/ZINB NLL ----
 * replaces DecoderSCVI.forward's heads (src/scvi/nn/_base_components.py:436-441:
 * px_scale = softmax(Linear(h)), px_dropout = Linear(h), px_rate = exp(library) * px_scale),
 * `px_r = exp(self.px_r)` (src/scvi/module/_vae.py:516-518), the distribution's log_prob
 * (log_zinb_positive / log_nb_positive) and `-log_prob(x).sum(-1)` (_vae.py:571).
 * The [B,G] tensors px_scale / px_rate / px_dropout / log_prob are never written to HBM.
 *   h [B,H]; Ws, Wp [G,H] (scale / dropout heads; Wp, bp NULL for NB); px_r [G];
 *   x rows via row_idx as in scvi_linear_fwd; library [B].
 * Outputs: reconst [B]; row_stats [B,4] = (row max of scale logits, sum exp(logit-max),
 *   exp(library)/sumexp, S = sum_g mu * d(reconst)/d(mu)) consumed by the backward.
 * Training fusion: when `dh` [B,H] is non-NULL the forward also emits
 *   dh[b,:] = g_reconst[b] * d(reconst[b])/d(h[b,:])   (g_reconst NULL = all ones),
 * computed in the same pass as the likelihood (the per-element derivatives share every
 * sub-expression with the log-probability), so the backward only has to produce the weight
 * gradients (call scvi_decoder_nll_bwd with dh = NULL).
 * SCVI_PREC_TC contract: library[b] >= log(max_g x[b, g]) - log 2 (true for the observed library log sum_g x):
 * the fp16 operand scale of a cell is derived from exp(library[b]).
 * precision SCVI_PREC_FP32: FFMA tiles, ws may be NULL.  SCVI_PREC_TC: tcgen05 tensor cores with an
 *   error-compensated fp16 split (fp32-level accuracy; needs H == 128) and a 256-byte aligned
 *   workspace of scvi_decoder_nll_fwd_ws_bytes() bytes. */
int64_t scvi_decoder_nll_fwd_ws_bytes(int64_t B, int64_t H, int64_t G, int likelihood, int precision);
int scvi_decoder_nll_fwd(const float* h, int64_t ldh, int64_t B, int64_t H, int64_t G, const float* Ws,
                         const float* bs, const float* Wp, const float* bp, const float* px_r,
                         const void* x, int64_t ldx, const int64_t* row_idx, const float* library,
                         int likelihood, int precision, float* reconst, float* row_stats,
                         const float* g_reconst, float* dh, int64_t lddh, void* ws, void* stream);
/* The parameter-only part of scvi_decoder_nll_fwd (SCVI_PREC_TC: fp16 hi/lo weight tiles, per-gene constants,
 * lgamma / digamma difference tables) written into the forward's workspace `ws` (same B, H, G, likelihood;
 * want_dh = the forward will be called with dh != NULL).  It depends on the weights only, so a caller can run
 * it on a second stream while the encoder computes and then call scvi_decoder_nll_fwd with
 * precision | SCVI_PREC_PREPACKED.  No-op for SCVI_PREC_FP32. */
int scvi_decoder_nll_prepack(int64_t B, int64_t H, int64_t G, const float* Ws, const float* bs, const float* Wp,
                             const float* bp, const float* px_r, int likelihood, int precision, int want_dh,
                             void* ws, void* stream);
/* g_reconst [B] = d(loss)/d(reconst).  Writes dh [B,H], dWs/dWp [G,H], dbs/dbp [G], dpx_r [G].
 * ws: scvi_decoder_nll_bwd_ws_bytes().  fwd_ws (nullable; SCVI_PREC_TC only): the workspace of the
 * scvi_decoder_nll_fwd call this backward belongs to (same B, G, likelihood, weights), still intact --
 * the packed weight tiles in it are reused instead of being rebuilt. */
int64_t scvi_decoder_nll_bwd_ws_bytes(int64_t B, int64_t H, int64_t G, int likelihood);
int scvi_decoder_nll_bwd(const float* h, int64_t ldh, int64_t B, int64_t H, int64_t G, const float* Ws,
                         const float* bs, const float* Wp, const float* bp, const float* px_r,
                         const void* x, int64_t ldx, const int64_t* row_idx, const float* library,
                         const float* row_stats, const float* g_reconst, int likelihood, int precision,
                         float* dh, int64_t lddh, float* dWs, float* dbs, float* dWp, float* dbp,
                         float* dpx_r, void* ws, const void* fwd_ws, void* stream);
/* materialising variant for inference-time getters (px.scale / px.mu / px.zi_logits):
 * any of the three outputs may be NULL. */
int scvi_decoder_params_fwd(const float* h, int64_t ldh, int64_t B, int64_t H, int64_t G, const float* Ws,
                            const float* bs, const float* Wp, const float* bp, const float* library,
                            float* px_scale, float* px_rate, float* px_dropout, int64_t ldo, void* stream);

/* ---- f1 / f2: per-row reductions of the semi-supervised and multi-modal losses ---------
 * out[r] = sum_d Normal(loc, scale).log_prob(value)[r, d] (torch.distributions.Normal.log_prob: -(v-m)^2 / (2 var)
 * - log(scale) - log(sqrt(2 pi))), as SCANVAE.loss uses it for the z1 terms (src/scvi/module/_scanvae.py:178-278).
 * `second` is the scale, or the variance when is_var != 0.  value / loc / second: contiguous [R, D]. */
int scvi_normal_logprob_sum_fwd(const float* value, const float* loc, const float* second, int64_t R, int64_t D,
                                int is_var, float* out, void* stream);
/* gradients times gout[r]; dvalue / dloc / dsecond [R, D] are each nullable. */
int scvi_normal_logprob_sum_bwd(const float* value, const float* loc, const float* second, const float* gout,
                                int64_t R, int64_t D, int is_var, float* dvalue, float* dloc, float* dsecond,
                                void* stream);
/* kl[r] = KL(Categorical(probs=p[r]) || Categorical(probs=q[r])) exactly as torch builds it (row-normalised probs,
 * logits = log(clamp(probs, eps, 1-eps)), inf where q == 0, 0 where p == 0): the label-prior term of SCANVAE.loss.
 * q: [R, C] (ldq = C) or one row shared by all (ldq = 0).  dp (nullable, [R, C]) = gout[r] * d kl[r] / d p[r, :];
 * kl may be NULL when only the gradient is wanted. */
int scvi_categorical_kl(const float* p, const float* q, int64_t ldq, int64_t R, int64_t C, const float* gout,
                        float* kl, float* dp, void* stream);

/* ---- a12: multi-tensor Adam over one flat arena ---------------------------------------
 * replaces torch.optim.Adam(lr=1e-3, eps=0.01, weight_decay=1e-6) built by
 * TrainingPlan.configure_optimizers (src/scvi/train/_trainingplans.py:475-482,497-519):
 * coupled L2, bias correction, same op order as torch's single-tensor path.
 * `step` is the 1-based step count; grad_scale multiplies g first (1/world for the DDP mean). */
int scvi_adam_step(float* p, const float* g, float* m, float* v, int64_t n, int64_t step, float lr,
                   float beta1, float beta2, float eps, float weight_decay, float grad_scale,
                   void* stream);
/* CUDA-graph friendly variant: the 1-based step is `step_counter[0] + 1` read on the device; the
 * last block of the update increments it, so a captured graph replays with the right bias
 * correction every time.  step_counter is int64[2]: {steps done, arrival ticket (keep 0)}. */
int scvi_adam_step_graph(float* p, const float* g, float* m, float* v, int64_t n, int64_t* step_counter,
                         float lr, float beta1, float beta2, float eps, float weight_decay,
                         float grad_scale, void* stream);
/* Gradient tensors -> the flat gradient arena of scvi_adam_step, one launch per 64 tensors: dst[offs[i] : offs[i]
 * + ns[i]] = src[i][0 : ns[i]] (fp32, contiguous; src[i] == NULL zero-fills the slot).  src / offs / ns are HOST
 * arrays; the table is passed by value to the kernel, so the call can be captured into a CUDA graph.  Stands where
 * torch's AccumulateGrad would add every gradient into a pre-zeroed `.grad` (one launch per parameter). */
int scvi_gather_segments(const void* const* src, const int64_t* offs, const int64_t* ns, int count, float* dst,
                         void* stream);
/* loss pieces of VAE.loss (src/scvi/module/_vae.py:573-580):
 * out[0] = mean_b(reconst + kl_weight * kl), out[1] = sum reconst, out[2] = sum kl.
 * kl_weight_dev (nullable, device float[1]) overrides `kl_weight` (graph replay: the KL warm-up
 * weight of src/scvi/train/_trainingplans.py:521-532 changes per epoch without re-capture). */
int scvi_elbo_reduce(const float* reconst, const float* kl, int64_t B, float kl_weight,
                     const float* kl_weight_dev, float* out3, void* stream);
/* ---- step prologue: everything the minibatch step needs that does not depend on the counts ------
 * One launch replacing, per step: the nn.Dropout mask of the encoder hidden layer and the rsample
 * noise (src/scvi/nn/_base_components.py:95-104,286-289), the one-hot batch covariate
 * (cov_cols[b] = batch_cols[idx[b]], src/scvi/nn/_base_components.py:139-169), the backward seeds
 * of the ELBO mean (as scvi_elbo_seed) and the BatchNorm num_batches_tracked increments.
 * mask [B,H] bytes (1 = keep, P(keep) = 1 - p_drop), noise [B,L] ~ N(0,1): Philox4x32-10 keyed by
 * `seed`, counter = (element, *step_counter, stream_id) -- the device-side step makes a captured CUDA
 * graph draw fresh numbers at every replay; stream_id (< 65536) separates data-parallel ranks.  These
 * are this library's own random streams (torch's generator is not reproduced).  Every output pointer is
 * nullable (skipped). */
int scvi_step_prologue(uint64_t seed, const int64_t* step_counter, int64_t stream_id, int64_t B, int64_t H,
                       int64_t L, float p_drop, uint8_t* mask, float* noise, const int64_t* idx,
                       const int32_t* batch_cols, int32_t* cov_cols, float kl_weight,
                       const float* kl_weight_dev, float scale, float* g_reconst, float* g_kl,
                       int64_t* nbt1, int64_t* nbt2, void* stream);
/* backward seeds of that mean: g_reconst[b] = scale / B, g_kl[b] = scale * kl_weight / B */
int scvi_elbo_seed(int64_t B, float kl_weight, const float* kl_weight_dev, float scale, float* g_reconst,
                   float* g_kl, void* stream);


/* Device address of page-locked (pinned, mapped) host memory: lets scvi_gather_rows_u8/_u16/scvi_gather_rows read a
 * count matrix that stays in HOST memory row by row over PCIe -- the minibatch gather of
 * AnnTorchDataset.__getitem__ (src/scvi/data/_anntorchdataset.py:143-147) done by the GPU, no host-side copy. */
int scvi_host_device_pointer(const void* host_ptr, void** dev_ptr);

/* ---- a13: data-parallel gradient mean fused with the optimizer, over NVLink peer memory -------------------
 * replaces torch DDP's gradient all-reduce (mean over replicas) followed by Adam.step under
 * `Trainer(strategy="ddp")` (reference tests/model/test_multigpu.py:29-36, src/scvi/model/_utils.py:25-34,
 * src/scvi/train/_trainingplans.py:475-519).  Single node: every rank allocates one exchange segment
 * (scvi_comm_alloc: cudaMalloc + zero fill + cudaIpcGetMemHandle, handle = 64 opaque bytes to ship to the other
 * ranks), opens the others' (scvi_comm_open, peer access enabled lazily) and lays it out as
 *     [gradient arena, step parity 0][gradient arena, step parity 1] ... [control block, scvi_comm_ctrl_bytes()]
 * scvi_allreduce_adam(slot): cross-GPU arrival barrier on the slot's flags (bounded wait: a missing peer traps),
 * g = sum_r peer_g[r][begin:end] in rank order (bit-identical on every rank), then the Adam update of
 * p/m/v[begin:end] with grad_scale 1/world, exactly scvi_adam_step_graph's arithmetic.  peer_g / peer_ctrl are HOST
 * arrays of `world` device pointers (this rank's own buffers at [rank]); every rank must issue the same sequence of
 * calls per slot.  The caller alternates the gradient buffer (parity) between consecutive steps.  advance_counter:
 * bump step_counter[0] when done (last bucket of the step).  g_sum (nullable): also store the reduced slice.
 * blocks: grid size (0 = default). */
int64_t scvi_comm_ctrl_bytes(void);
int scvi_comm_alloc(int64_t bytes, void** dev_ptr, void* ipc_handle64);
int scvi_comm_open(const void* ipc_handle64, void** dev_ptr);
int scvi_comm_close(void* dev_ptr);
int scvi_comm_free(void* dev_ptr);
int scvi_allreduce_adam(const void* const* peer_g, void* const* peer_ctrl, int world, int rank, int slot,
                        float* p, float* m, float* v, float* g_sum, int64_t begin, int64_t end,
                        int64_t* step_counter, int advance_counter, float lr, float beta1, float beta2,
                        float eps, float weight_decay, int blocks, void* stream);
/* Sharded form of the same exchange (one launch at the end of the step): rank r reduces the r-th 1/world slice of
 * the gradient arena (rank order), applies Adam to it with its own moments and stores the updated PARAMETERS into
 * every rank's parameter arena peer_p[*] (the parameter arenas live in the exchange segments too); arrival barrier
 * on flags `slot`, completion barrier on flags `slot + 1`: when the call's kernel ends, peer_p[rank] holds every
 * slice.  (N-1)/N of the arena read and written per rank instead of N-1 arenas read; replicas bit-identical by
 * construction.  m / v are only maintained for the rank's own slice.  n = arena floats (multiple of 4). */
int scvi_allreduce_adam_sharded(const void* const* peer_g, void* const* peer_p, void* const* peer_ctrl, int world,
                                int rank, int slot, float* m, float* v, int64_t n, int64_t* step_counter, float lr,
                                float beta1, float beta2, float eps, float weight_decay, int blocks, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SCVI_B200_H */
