/*
 * pclip.h -- C ABI of libpclip.so, the B200 (sm_100a) implementation of protein-clip's contrastive hot path.
 *
 * The reference (pan-emily/protein-clip) has NO native / FFI boundary: its hot path is a handful of
 * Python functions built from torch library ops.  Each entry point below therefore cites the Python
 * call site it replaces ("ref:" = path relative to the reference checkout).  The Python host code in
 * protein-clip_b200/pclip binds these with ctypes (see INTEGRATION.md for the exact stubs) and keeps
 * the reference's module-level names call-compatible.
 *
 * Conventions
 *   - every function returns 0 (PCLIP_OK) or a negative pclip_status; pclip_last_error() gives the
 *     thread-local message of the last failure.
 *   - all data pointers are DEVICE pointers unless the comment says HOST.
 *   - nothing here allocates or frees caller memory; scratch comes from a caller workspace whose size
 *     is given by the matching *_workspace_bytes() query (1024-byte aligned).
 *   - `stream` is a cudaStream_t passed as void*; all work is enqueued on it, nothing synchronises.
 *   - no global state; calls are re-entrant per (stream, workspace).
 *   - there is NO CPU fallback: without a CUDA device every compute entry point fails with
 *     PCLIP_ERR_CUDA.
 */
#ifndef PCLIP_H_
#define PCLIP_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* pclip_stream_t; /* cudaStream_t */

enum pclip_status {
  PCLIP_OK = 0,
  PCLIP_ERR_INVALID = -1,     /* bad argument (null pointer, negative size, misaligned pointer) */
  PCLIP_ERR_UNSUPPORTED = -2, /* shape / dtype combination with no kernel (e.g. D != 128 in bf16 mode) */
  PCLIP_ERR_CUDA = -3,        /* a CUDA runtime / driver call failed; message has the CUDA error string */
  PCLIP_ERR_WORKSPACE = -4    /* workspace too small or misaligned (needs 1024 B alignment) */
};

enum pclip_dtype { PCLIP_F32 = 0, PCLIP_BF16 = 1 };

/* arithmetic the similarity contractions run in */
enum pclip_compute {
  PCLIP_COMPUTE_F32 = 0, /* FP32 FFMA on CUDA cores: matches the reference's fp32 path to ~1e-6 */
  PCLIP_COMPUTE_BF16 = 1 /* BF16 operands, FP32 accumulate, tcgen05 tensor cores (TMA + TMEM); D == 128 */
};

/* A strided 2-D row block living on the device. Lists of these express GradCache's "list of leaves"
 * without a torch.cat (ref: GradCache cat_input_tensor, modules/training_utils.py:247-249), and a
 * `.t()` view (ref: modules/training_utils.py:69) through its strides. */
typedef struct {
  const void* ptr;    /* device pointer to element (0,0) */
  int64_t rows;       /* number of rows in this chunk */
  int64_t row_stride; /* in elements */
  int64_t col_stride; /* in elements */
} pclip_chunk_t;

int pclip_version(void);
const char* pclip_last_error(void);
/* number of kernels launched by this thread through the library since the last reset (bench bookkeeping) */
int64_t pclip_launch_count(void);
void pclip_launch_count_reset(void);

/* ------------------------------------------------------------------------------------------------
 * (1) Encoder tail: L2-normalise + temperature scale.
 *     ref: modules/models.py:50-51   out = F.normalize(e, dim=-1) * exp(tau / 2)
 *     eps = 1e-12 -> x / max(||x||, eps).  With eps == 0 and tau == NULL this is the FILIP token
 *     normalisation, ref: modules/models.py:238-239 (h / h.norm(dim=-1, keepdim=True), no eps).
 * ---------------------------------------------------------------------------------------------- */
int pclip_normalize_scale_fwd(const float* e, int64_t rows, int32_t D, const float* tau /* dev [1] or NULL */,
                              float eps, float* out /* [rows, D] */, float* inv_norm /* [rows] saved */,
                              pclip_stream_t stream);
/* scratch of the grid-wide sums below (dtau, row-wise loss): summed in a fixed order, so results are reproducible
 * run to run (no float atomics) */
size_t pclip_rowsum_workspace_bytes(int64_t rows);
/* de = (s / n) * (g - nhat * <g, nhat>),  dtau += 0.5 * sum(g * out)   (dtau may be NULL: then no workspace is needed) */
int pclip_normalize_scale_bwd(const float* grad_out, const float* out, const float* inv_norm, int64_t rows,
                              int32_t D, const float* tau, float* de, float* dtau /* dev [1], accumulated */,
                              void* workspace, size_t workspace_bytes, pclip_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * (2) CLIP InfoNCE, symmetric cross entropy over S = X Y^T.
 *     ref: modules/training_utils.py:74-89  (_contrastive_loss)
 *          modules/training_utils.py:247-265 (_contrastive_loss_gc: same math on concatenated leaves)
 *     loss = 1/(2B) * sum_i [ log R_i + log C_i - 2 S_ii ],  R_i = sum_j exp S_ij, C_j = sum_i exp S_ij
 *     dS   = 1/(2B) * [ E/R_i + E/C_j - 2 I ],  dX = dS Y,  dY = dS^T X
 *     exp is the reference's plain exp (no max subtraction): results are identical wherever the
 *     reference's are finite.
 *
 *     Every contraction is expressed as a "side": rows of `a` against all rows of `b`.  R comes from the
 *     side (X, Y), C from the side (Y, X); dX from (X, Y, 1/R, 1/C), dY from (Y, X, 1/C, 1/R).  A rank of a
 *     row-sharded job passes its own rows as `a` and the all-gathered other modality as `b`, so that no
 *     gradient has to be exchanged.
 * ---------------------------------------------------------------------------------------------- */

/* pack chunk lists into a contiguous [rows, D] matrix of `out_dtype`; optionally also the transposed copy
 * outT [D, ldT] (ldT >= rows, multiple of 64; needed by the bf16 gradient kernel).  NULL = not wanted. */
int pclip_pack(const pclip_chunk_t* chunks /* HOST */, int32_t n, int32_t in_dtype, int64_t rows, int32_t D,
               int32_t out_dtype, void* out, void* outT, int64_t ldT, pclip_stream_t stream);
/* both modalities of a step (two chunk lists of the same total shape) with ONE launch, no transposed copies
 * (ref: the torch.cat of GradCache's two leaf lists, modules/training_utils.py:247-249) */
int pclip_pack_pair(const pclip_chunk_t* x_chunks /* HOST */, int32_t nx, const pclip_chunk_t* y_chunks /* HOST */,
                    int32_t ny, int32_t in_dtype, int64_t rows, int32_t D, int32_t out_dtype, void* out_x, void* out_y,
                    pclip_stream_t stream);
/* outT[d, r] = in[r, d] for packed bf16 [rows, D] (used after an all-gather) */
int pclip_transpose_bf16(const void* in, int64_t rows, int32_t D, void* outT, int64_t ldT, pclip_stream_t stream);

typedef struct {
  const void* a;       /* [Ma, D] packed, compute dtype */
  const void* b;       /* [Nb, D] packed, compute dtype */
  int64_t Ma, Nb;
  int64_t diag_offset; /* row i of `a` is the positive of row diag_offset + i of `b` */
  float* rowsum;       /* out [Ma]: sum_j exp(a_i . b_j) */
  float* diag;         /* out [Ma]: a_i . b_(diag_offset+i), or NULL */
} pclip_clip_fwd_side_t;

size_t pclip_clip_rowsumexp_workspace_bytes(int64_t Ma, int64_t Nb, int32_t D, int32_t compute);
int pclip_clip_rowsumexp(const pclip_clip_fwd_side_t* sides /* HOST */, int32_t n_sides /* 1 or 2 */, int32_t D,
                         int32_t compute, void* workspace, size_t workspace_bytes, pclip_stream_t stream);

/* Single-pass variant (bf16 compute only): one sweep over S = a b^T gives
 *   rowsum[i] = sum_j exp(a_i . b_j)                 (i < Ma; complete)
 *   colsum[j] = sum_{i < Ma} exp(a_i . b_j)          (j < Nb; partial over THIS call's rows: a row-sharded job
 *                                                     all-reduces it)
 *   diag[i]   = a_i . b_(diag_offset + i)
 * ref: modules/training_utils.py:85-88 (exp_logits.sum(1) and exp_logits.sum(0) of ONE logits matrix). */
size_t pclip_clip_sumexp_workspace_bytes(int64_t Ma, int64_t Nb, int32_t D, int32_t compute);
int pclip_clip_sumexp(const void* a /* [Ma,D] bf16 */, const void* b /* [Nb,D] bf16 */, int64_t Ma, int64_t Nb,
                      int64_t diag_offset, int32_t D, int32_t compute, float* rowsum /* [Ma] */,
                      float* colsum /* [Nb] */, float* diag /* [Ma] */, void* workspace, size_t workspace_bytes,
                      pclip_stream_t stream);

/* per-row loss terms -> row_lse/col_lse, 1/R, 1/C and loss[0] (+)= scale * sum_i(log R_i + log C_i - 2 diag_i) */
int pclip_clip_finalize(const float* rowsum, const float* colsum, const float* diag, int64_t M, float scale,
                        float* row_lse, float* col_lse, float* rinv, float* cinv, float* loss /* [1] */,
                        pclip_stream_t stream);

/* Row-sharded job, after the all-gather of every rank's forward record
 * stats_all [P][Nb + 2 Ma] = [colsum partial (Nb) | rowsum (Ma) | diag (Ma)]: column sums added in rank order, R and S_ii
 * taken by global row, then the same per-row terms and loss as pclip_clip_finalize over all Nb rows, in ONE launch; the
 * loss terms are added in a fixed order (bit-reproducible; P = 1 is the single-GPU case and gives the same bits). */
size_t pclip_clip_finalize_sharded_workspace_bytes(int64_t Nb);
int pclip_clip_finalize_sharded(const float* stats_all, int32_t P, int64_t Nb, int64_t Ma, float scale, float* row_lse,
                                float* col_lse, float* rinv, float* cinv, float* loss /* [1] */, void* workspace,
                                size_t workspace_bytes, pclip_stream_t stream);

typedef struct {
  const void* a;     /* [Ma, D] */
  const void* b;     /* [Nb, D] */
  const void* bT;    /* [D, ldbT] transposed copy of b (bf16 compute only; NULL for fp32) */
  int64_t ldbT;
  int64_t Ma, Nb;
  int64_t diag_offset;
  const float* rinv; /* [Ma] 1 / (sum-exp of row i) */
  const float* cinv; /* [Nb] 1 / (sum-exp of column j) */
  float* da;         /* out [Ma, D] fp32: scale * (sum_j exp(s_ij) (rinv_i + cinv_j) b_j - 2 b_(diag_offset+i)) */
} pclip_clip_bwd_side_t;

size_t pclip_clip_grad_workspace_bytes(int64_t Ma, int64_t Nb, int32_t D, int32_t compute);
/* `scale` is 1/(2B) times autograd's grad_output when that is known on the host; gscale (dev [1], may be
 * NULL) multiplies it on the device (GradScaler passes 65536, ref: modules/training_utils.py:139). */
int pclip_clip_grad(const pclip_clip_bwd_side_t* sides /* HOST */, int32_t n_sides, int32_t D, int32_t compute,
                    float scale, const float* gscale, void* workspace, size_t workspace_bytes,
                    pclip_stream_t stream);

/* Single-GPU composite: pack -> row/column sum-exp -> loss (-> gradients for grad_output == 1).
 * x / y are lists of row chunks (n == 1 for a plain tensor; y may be a transposed view through strides). */
size_t pclip_clip_workspace_bytes(int64_t B, int32_t D, int32_t compute);
int pclip_clip_loss(const pclip_chunk_t* x_chunks /* HOST */, int32_t nx, const pclip_chunk_t* y_chunks /* HOST */,
                    int32_t ny, int32_t in_dtype, int64_t B, int32_t D, int32_t compute,
                    float* loss /* [1] */, float* row_lse /* [B] or NULL */, float* col_lse /* [B] or NULL */,
                    float* dx /* [B, D] fp32 or NULL */, float* dy, void* workspace, size_t workspace_bytes,
                    pclip_stream_t stream);

/* Small-batch CLIP step in ONE launch (B <= pclip_clip_small_max_batch(), D == 128; fp32 arithmetic): loss and -- when
 * dx_chunks / dy_chunks are given -- the gradient of every leaf, multiplied by 1/(2B) and by *gscale (GradScaler's
 * scale, ref: modules/training_utils.py:139; NULL = 1), written straight into the caller's per-leaf fp32 buffers
 * (chunk i of dx has the rows of chunk i of x, col_stride 1).  This is the GradCache configuration
 * (ref: modules/training_utils.py:247-265 on 16 + 16 leaves of [16, 128]): no torch.cat, no pack, no logits in HBM. */
int32_t pclip_clip_small_max_batch(void);
size_t pclip_clip_small_workspace_bytes(int64_t B);
int pclip_clip_small_fused(const pclip_chunk_t* x_chunks /* HOST */, int32_t nx, const pclip_chunk_t* y_chunks /* HOST */,
                           int32_t ny, int32_t in_dtype, int64_t B, int32_t D, const float* gscale /* dev [1] or NULL */,
                           float* loss /* [1] */, const pclip_chunk_t* dx_chunks /* HOST, or NULL */,
                           const pclip_chunk_t* dy_chunks /* HOST, or NULL */, void* workspace, size_t workspace_bytes,
                           pclip_stream_t stream);

/* out[i] = in[i] * (*g)   -- applies autograd's scalar grad_output on the device, no host sync */
int pclip_scale_by_scalar(const float* in, const float* g /* dev [1] */, int64_t n, float* out,
                          pclip_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * (3) Similarity scores / rank-of-positive / top-k retrieval.
 *     ref: modules/training_utils.py:157-172 (get_logits: exp(x y^T))
 *          modules/visualizations.py:38-52   (cosine matrix: x y^T * exp(-tau))
 *          modules/training_utils.py:219-222 (rank = #{j: s_j > s_pos} + 1, strict >)
 *          topk.py:47-61                     (top-k accuracy = cumulative rank histogram)
 * ---------------------------------------------------------------------------------------------- */
enum pclip_score_op { PCLIP_SCORE_DOT = 0, PCLIP_SCORE_EXP = 1 };

/* out[q, n] = op(q . bank_n) * (neg_tau ? exp(-*neg_tau) : 1);  q, bank packed fp32 */
int pclip_scores(const float* q, const float* bank, int64_t Q, int64_t N, int32_t D, int32_t op,
                 const float* neg_tau /* dev [1] or NULL */, float* out, pclip_stream_t stream);

/* rank[q] = 1 + #{ n : q.bank_n > q.bank_pos[q] }  in fp32 (every score computed by the same FFMA chain).
 * One call ranks ALL Q queries against the bank (topk.py:47-51 loops 256 single-query calls). */
size_t pclip_rank_workspace_bytes(int64_t Q);
int pclip_rank_of_positive(const float* q, const float* bank, const int64_t* pos, int64_t Q, int64_t N, int32_t D,
                           int32_t* rank, void* workspace, size_t workspace_bytes, pclip_stream_t stream);

size_t pclip_topk_workspace_bytes(int64_t Q, int64_t N, int32_t D, int32_t k, int32_t compute);
/* top-k by dot product, descending, ties -> lower index.  idx [Q,k] int64 (+ index_base), score [Q,k] fp32.
 * compute == F32: exact fp32 scores (one FFMA chain per score, d ascending).
 * compute == BF16: tensor-core candidate selection (k + margin candidates per query) followed by fp32
 * re-scoring of the candidates with the same FFMA chain; `n_unsafe` (dev [1], may be NULL) counts queries
 * whose candidate margin could not prove exactness -- those rows are recomputed by the fp32 kernel. */
int pclip_topk(const float* q, const float* bank, int64_t Q, int64_t N, int32_t D, int32_t k, int32_t compute,
               int64_t index_base, int64_t* idx, float* score, int32_t* n_unsafe, void* workspace,
               size_t workspace_bytes, pclip_stream_t stream);
/* Persistent bf16 image of a bank for repeated bf16 top-k calls (an embedding bank is built once and queried many
 * times): pclip_topk converts the fp32 bank on every call (3 N D bytes of HBM traffic), which
 * dominates when a call carries few queries.  `prepared` is caller memory of pclip_topk_bank_bytes(N, D) bytes,
 * 1024-byte aligned; `bank` (fp32) is still needed by pclip_topk_prepared for the exact re-scoring. */
size_t pclip_topk_bank_bytes(int64_t N, int32_t D);
int pclip_topk_prepare_bank(const float* bank, int64_t N, int32_t D, void* prepared, size_t prepared_bytes,
                            pclip_stream_t stream);
int pclip_topk_prepared(const float* q, const float* bank, const void* prepared, int64_t Q, int64_t N, int32_t D,
                        int32_t k, int64_t index_base, int64_t* idx, float* score, int32_t* n_unsafe, void* workspace,
                        size_t workspace_bytes, pclip_stream_t stream);

/* merge per-shard (idx, score) candidate lists [n_lists, Q, k] into the global top-k (bank-sharded mode) */
int pclip_topk_merge(const int64_t* idx_in, const float* score_in, int64_t Q, int32_t n_lists, int32_t k,
                     int64_t* idx, float* score, pclip_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * (4) FILIP token-wise late interaction.
 *     ref: modules/models.py:292-356 (_filip_similarity_score), :247-264 (_masked_mean)
 *          modules/training_utils_filip.py:4-39 (_contrastive_loss row-wise on BOTH matrices, _compute_loss)
 *     S[a,b,t,s] = <hA[a,t], hB[b,s]> / tau over the valid prefixes lenA[a] x lenB[b] (right-padded masks);
 *     simA[a,b] = sum_{t < lenA[a]} max_{s < lenB[b]} S / max(lenA[a], 1e-6)
 *     simB[a,b] = sum_{s < lenB[b]} max_{t < lenA[a]} S / max(lenB[b], 1e-6)
 *     argA[a,b,t] / argB[b,a,s] (NOTE the [Bb,Ba,TB] layout of argB) record the maximising token, lowest index
 *     on exact ties, for the backward pass.  hA / hB are what _filip_similarity_score receives (already
 *     normalised by the caller, ref: models.py:238-239 -> pclip_normalize_scale_fwd with eps = 0, tau = NULL).
 *     tau must be > 0 (the reference's nn.Parameter(1.0)); tau <= 0 yields NaN outputs.
 *     Ba != Bb is allowed: a rank of an A-row-sharded job passes its rows of hA and all of hB.
 *     compute == BF16 needs D == 128 and TA, TB <= 32767.
 * ---------------------------------------------------------------------------------------------- */
size_t pclip_filip_workspace_bytes(int64_t Ba, int64_t Bb, int64_t TA, int64_t TB, int32_t D, int32_t compute);

int pclip_filip_fwd(const float* hA /* [Ba,TA,D] */, const float* hB /* [Bb,TB,D] */,
                    const int32_t* lenA /* [Ba] */, const int32_t* lenB /* [Bb] */, int64_t Ba, int64_t Bb,
                    int64_t TA, int64_t TB, int32_t D, const float* tau /* dev [1] */, int32_t compute,
                    float* simA /* [Ba,Bb] */, float* simB /* [Ba,Bb] */, int16_t* argA /* [Ba,Bb,TA] */,
                    int16_t* argB /* [Bb,Ba,TB] */, void* workspace, size_t workspace_bytes,
                    pclip_stream_t stream);

/* gA / gB: upstream gradients of simA / simB [Ba,Bb].  dhA [Ba,TA,D], dhB [Bb,TB,D] are OVERWRITTEN,
 * dtau [1] is accumulated (dtau += -(sum gA simA + sum gB simB) / tau). */
int pclip_filip_bwd(const float* hA, const float* hB, const int32_t* lenA, const int32_t* lenB, int64_t Ba,
                    int64_t Bb, int64_t TA, int64_t TB, int32_t D, const float* tau, int32_t compute,
                    const float* simA, const float* simB, const int16_t* argA, const int16_t* argB,
                    const float* gA, const float* gB, float* dhA, float* dhB, float* dtau, void* workspace,
                    size_t workspace_bytes, pclip_stream_t stream);

/* row-wise InfoNCE on a [B, Bc] logits matrix whose positives sit at column diag_offset + i:
 *   loss (+)= scale * sum_i (log sum_j exp L_ij - L_i,diag);  dlogits = scale * (softmax_row - onehot)
 * ref: modules/training_utils_filip.py:15-23 */
int pclip_rowwise_infonce(const float* logits, int64_t B, int64_t Bc, int64_t diag_offset, float scale,
                          float* loss /* [1], accumulated */, float* dlogits /* [B,Bc] or NULL */,
                          void* workspace /* pclip_rowsum_workspace_bytes(B) */, size_t workspace_bytes,
                          pclip_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * Fused encoder head, forward (ref: modules/models.py:46-51 Encoder.forward after the ESM trunk, and
 * :192-196 FILIPEncoder.forward): projection Din -> E, per-residue FFN, then
 *   out_tokens != NULL : the per-residue embeddings [Bs, T, E]                               (FILIPEncoder)
 *   out_pooled != NULL : masked mean over residues (count clamped to >= 1, :75-90), pooled FFN, F.normalize(eps),
 *                        * exp(tau / 2) -> [Bs, E]                                           (Encoder)
 * Inference only: no dropout, nothing kept for a backward.  fp32, E = 128, Din % 4 == 0.  An FFN of depth h is h - 1
 * layers with a LayerNorm (Linear, ReLU, LayerNorm: ln_weight / ln_bias set) followed by one plain Linear (both NULL),
 * i.e. the nn.Sequential of models.py:66-73 in order.  All parameter pointers 16-byte aligned.
 * ---------------------------------------------------------------------------------------------- */
typedef struct {
  const float* weight;    /* [E, K] row-major (nn.Linear.weight) */
  const float* bias;      /* [E] */
  const float* ln_weight; /* [E] LayerNorm weight, or NULL: no ReLU + LayerNorm after this Linear */
  const float* ln_bias;   /* [E] */
} pclip_linear_t;
size_t pclip_encoder_head_workspace_bytes(int64_t Bs, int64_t T);   /* pooled mode only */
int pclip_encoder_head_fwd(const float* hidden /* [Bs,T,Din] */, const int64_t* mask /* [Bs,T] 0/1 */, int64_t Bs, int64_t T,
                           int32_t Din, int32_t E, const pclip_linear_t* projection /* HOST, 1 */,
                           const pclip_linear_t* token_ffn /* HOST */, int32_t n_token,
                           const pclip_linear_t* pooled_ffn /* HOST */, int32_t n_pooled, const float* tau /* [1] or NULL */,
                           float ln_eps, float norm_eps, float* out_tokens, float* out_pooled, void* workspace,
                           size_t workspace_bytes, pclip_stream_t stream);

/* pclip_rowwise_infonce for callers that cannot allocate or own a workspace (CUDA graph capture, serving loops):
 * same arguments, arithmetic, summation order and results, but the scratch of the ordered loss sum lives in a
 * file-scope static __device__ array of the library (16.6 KB, held by every GPU that loads the library until exit).
 * No counterpart in the reference.
 * Limit: B <= pclip_rowwise_infonce_noalloc_max_rows() = 32768, larger B returns PCLIP_ERR_UNSUPPORTED.  Calls no
 * cudaMalloc / cudaFree.  All calls on one device share the array: calls that overlap on the GPU (different streams,
 * no ordering between them) race; order them or use pclip_rowwise_infonce with one workspace per stream. */
int64_t pclip_rowwise_infonce_noalloc_max_rows(void);
int pclip_rowwise_infonce_noalloc(const float* logits, int64_t B, int64_t Bc, int64_t diag_offset, float scale,
                                  float* loss /* [1], accumulated */, float* dlogits /* [B,Bc] or NULL */,
                                  pclip_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* PCLIP_H_ */
