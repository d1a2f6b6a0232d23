// Similarity scores, rank-of-positive and exact fp32 top-k (PCLIP_COMPUTE_F32).
//   ref: modules/training_utils.py:157-172 (get_logits = exp(x y^T)), :219-222 (rank = #{j: s_j > s_pos} + 1),
//        modules/visualizations.py:51-52 (cosine = x y^T exp(-tau)), topk.py:47-61 (top-k accuracy from ranks).
// All scores come from simt_tile.cuh's FFMA chain, so "s_j > s_pos" compares numbers produced by the same
// arithmetic (a tie is a tie, as in the reference where both sides come from one matmul).
#include <float.h>

#include "common.cuh"
#include "simt_tile.cuh"

namespace pclip {

// ---------------------------------------------------------------------------------------------------------
// dense score matrix
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) scores_kernel(const float* __restrict__ q, const float* __restrict__ bank,
                                                     int64_t Q, int64_t N, int D, int op,
                                                     const float* __restrict__ neg_tau, float* __restrict__ out) {
  __shared__ __align__(16) float As[kBK][kBM + kPad];
  __shared__ __align__(16) float Bs[kBK][kBN + kPad];
  const int64_t row0 = (int64_t)blockIdx.y * kBM, col0 = (int64_t)blockIdx.x * kBN;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  float acc[4][4];
  score_tile(q, bank, Q, N, D, row0, col0, As, Bs, acc);
  const float mul = neg_tau ? expf(-neg_tau[0]) : 1.0f;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int64_t r = row0 + ty * 4 + i;
    if (r >= Q) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int64_t c = col0 + tx * 4 + j;
      if (c < N) out[r * N + c] = (op == PCLIP_SCORE_EXP ? expf(acc[i][j]) : acc[i][j]) * mul;
    }
  }
}

// ---------------------------------------------------------------------------------------------------------
// rank of the positive
// ---------------------------------------------------------------------------------------------------------
// pos_score[q] = q . bank[pos[q]] with the same ascending-k FFMA chain as score_tile; rank[q] = 1
__global__ void pos_score_kernel(const float* __restrict__ q, const float* __restrict__ bank,
                                 const int64_t* __restrict__ pos, int64_t Q, int D, float* __restrict__ pos_score,
                                 int32_t* __restrict__ rank) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= Q) return;
  const float* a = q + i * D;
  const float* b = bank + pos[i] * D;
  float acc = 0.f;
  for (int k = 0; k < D; ++k) acc = fmaf(a[k], b[k], acc);
  pos_score[i] = acc;
  rank[i] = 1;
}

__global__ void __launch_bounds__(256) rank_kernel(const float* __restrict__ q, const float* __restrict__ bank,
                                                   const float* __restrict__ pos_score, int64_t Q, int64_t N, int D,
                                                   int tiles_per_split, int32_t* __restrict__ rank) {
  __shared__ __align__(16) float As[kBK][kBM + kPad];
  __shared__ __align__(16) float Bs[kBK][kBN + kPad];
  const int64_t row0 = (int64_t)blockIdx.x * kBM;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  float ps[4];
  int cnt[4] = {0, 0, 0, 0};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int64_t r = row0 + ty * 4 + i;
    ps[i] = r < Q ? pos_score[r] : FLT_MAX;
  }
  const int64_t ntiles = ceil_div(N, kBN);
  const int64_t t0 = (int64_t)blockIdx.y * tiles_per_split;
  const int64_t t1 = t0 + tiles_per_split < ntiles ? t0 + tiles_per_split : ntiles;
  float acc[4][4];
  for (int64_t t = t0; t < t1; ++t) {
    const int64_t col0 = t * kBN;
    score_tile(q, bank, Q, N, D, row0, col0, As, Bs, acc);
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (col0 + tx * 4 + j < N && acc[i][j] > ps[i]) ++cnt[i];
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) cnt[i] += __shfl_xor_sync(0xffffffffu, cnt[i], o);
    const int64_t r = row0 + ty * 4 + i;
    if (tx == 0 && r < Q && cnt[i]) atomicAdd(&rank[r], cnt[i]);  // integer: order-independent
  }
}

// ---------------------------------------------------------------------------------------------------------
// exact top-k.  CTA = 64 queries; per bank tile the 64x64 scores go through shared memory and one thread per
// query scans its row in ascending bank index, inserting into a sorted list (score desc, index asc).  A later
// equal score never displaces an earlier one, so ties resolve to the lower index.
// ---------------------------------------------------------------------------------------------------------
constexpr int kMaxK = 64;
constexpr int kMaxLists = 296;   // most bank splits of the list-mode pass (a handful of rows: two CTAs per SM scan the bank)
// List mode: the number of listed rows is only known on the device, so the bank split is chosen there.  The partial
// lists ([split][listed row][k]) get a fixed budget of entries; a few listed rows use all kMaxLists splits, many rows
// fewer, longer ones.  (A host-side plan had to assume all Q rows could be listed: 68 splits at Q = 262 144, 18 ms for
// 3 rows.)
constexpr size_t kListPartialBudget = size_t(2) << 30;
static int64_t list_mode_entries(int64_t Q, int k) {
  const int64_t budget = (int64_t)(kListPartialBudget / (sizeof(int64_t) + sizeof(float)));
  return std::max<int64_t>(Q * k, std::min<int64_t>(budget, (int64_t)kMaxLists * Q * k));
}
__device__ __forceinline__ void list_mode_plan(int64_t count, int k, int64_t entries, int64_t ntiles, int max_splits,
                                               int* tps, int* n_used) {
  int64_t n = entries / (count * k);
  n = n < 1 ? 1 : (n > max_splits ? max_splits : n);
  n = n > ntiles ? ntiles : n;
  *tps = (int)((ntiles + n - 1) / n);
  *n_used = (int)((ntiles + *tps - 1) / *tps);
}

__global__ void __launch_bounds__(256) topk_kernel(const float* __restrict__ q, const float* __restrict__ bank,
                                                   int64_t Q, int64_t N, int D, int k, int tiles_per_split,
                                                   int64_t index_base, int64_t* __restrict__ idx_out,
                                                   float* __restrict__ score_out, const int32_t* __restrict__ qlist,
                                                   const int32_t* __restrict__ qcount, int64_t list_entries) {
  // list mode (qlist != NULL): logical row r is query qlist[r]; only the first *qcount logical rows exist.
  int64_t Qstride = Q;
  if (qcount) {
    Q = *qcount;
    if (Q <= 0) return;
    int n_used;
    list_mode_plan(Q, k, list_entries, ceil_div(N, kBN), (int)gridDim.y, &tiles_per_split, &n_used);
    if ((int)blockIdx.y >= n_used) return;
    Qstride = Q;
  }
  extern __shared__ __align__(16) uint8_t dyn[];
  float(*As)[kBM + kPad] = reinterpret_cast<float(*)[kBM + kPad]>(dyn);
  float(*Bs)[kBN + kPad] = As + kBK;
  float(*S)[kBN + 1] = reinterpret_cast<float(*)[kBN + 1]>(Bs + kBK);
  float* lscore = reinterpret_cast<float*>(S + kBM);      // [64][k]
  int64_t* lidx = reinterpret_cast<int64_t*>(lscore + kBM * kMaxK);  // [64][k]
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int64_t ntiles = ceil_div(N, kBN);
  const int64_t t0 = (int64_t)blockIdx.y * tiles_per_split;
  const int64_t t1 = t0 + tiles_per_split < ntiles ? t0 + tiles_per_split : ntiles;
  // Row blocks are strided over gridDim.x: in list mode the live row count is only known on the device, and a grid
  // sized for all Q rows meant ~10^6 blocks that exit at once (24 ms per call at Q = 262 144 for 3 listed queries).
  for (int64_t rblk = blockIdx.x; rblk * kBM < Q; rblk += gridDim.x) {
  const int64_t row0 = rblk * kBM;
  int filled = 0;  // valid for tid < 64 (row owner)
  float acc[4][4];
  for (int64_t t = t0; t < t1; ++t) {
    const int64_t col0 = t * kBN;
    score_tile(q, bank, Q, N, D, row0, col0, As, Bs, acc, qlist);
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) S[ty * 4 + i][tx * 4 + j] = acc[i][j];
    __syncthreads();
    if (tid < kBM && row0 + tid < Q) {
      float* ls = lscore + tid * kMaxK;
      int64_t* li = lidx + tid * kMaxK;
      const int ncol = (int)(N - col0 < kBN ? N - col0 : kBN);
      for (int c = 0; c < ncol; ++c) {
        const float s = S[tid][c];
        if (filled == k && !(s > ls[k - 1])) continue;
        int p = filled < k ? filled : k - 1;  // slot that is freed / appended
        while (p > 0 && ls[p - 1] < s) {      // strictly smaller entries move down; equal ones stay ahead
          ls[p] = ls[p - 1];
          li[p] = li[p - 1];
          --p;
        }
        ls[p] = s;
        li[p] = col0 + c;
        if (filled < k) ++filled;
      }
    }
    __syncthreads();
  }
  if (tid < kBM && row0 + tid < Q) {
    const int64_t r = row0 + tid;
    const int64_t o = ((int64_t)blockIdx.y * Qstride + r) * k;
    for (int p = 0; p < k; ++p) {
      const bool have = p < filled;
      idx_out[o + p] = have ? lidx[tid * kMaxK + p] + index_base : -1;
      score_out[o + p] = have ? lscore[tid * kMaxK + p] : -FLT_MAX;
    }
  }
  __syncthreads();
  }
}

// merge n_lists sorted candidate lists per query: lists laid out [n_lists][Q][k]; (score desc, index asc)
__global__ void topk_merge_kernel(const int64_t* __restrict__ idx_in, const float* __restrict__ score_in, int64_t Q,
                                  int n_lists, int k, int64_t* __restrict__ idx, float* __restrict__ score,
                                  const int32_t* __restrict__ qlist, const int32_t* __restrict__ qcount,
                                  int64_t list_entries, int64_t ntiles) {
  const int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (r >= (qcount ? (int64_t)*qcount : Q)) return;
  if (qcount) {   // list mode: the split the scan kernel derived from the live row count
    Q = *qcount;
    int tps;
    list_mode_plan(Q, k, list_entries, ntiles, n_lists, &tps, &n_lists);
  }
  const int64_t wr = qlist ? (int64_t)qlist[r] : r;  // where the merged row is written
  int head[kMaxLists];
  for (int l = 0; l < n_lists; ++l) head[l] = 0;
  for (int p = 0; p < k; ++p) {
    int best = -1;
    float bs = 0.f;
    int64_t bi = 0;
    for (int l = 0; l < n_lists; ++l) {
      if (head[l] >= k) continue;
      const int64_t o = ((int64_t)l * Q + r) * k + head[l];
      const int64_t ci = idx_in[o];
      if (ci < 0) continue;
      const float cs = score_in[o];
      if (best < 0 || cs > bs || (cs == bs && ci < bi)) {
        best = l;
        bs = cs;
        bi = ci;
      }
    }
    if (best < 0) {
      idx[wr * k + p] = -1;
      score[wr * k + p] = -FLT_MAX;
    } else {
      idx[wr * k + p] = bi;
      score[wr * k + p] = bs;
      ++head[best];
    }
  }
}

static void plan_n_split(int64_t Q, int64_t N, int max_lists, int* nsplit, int* tps) {
  const int64_t rb = ceil_div(Q, kBM), nt = ceil_div(N, kBN);
  int64_t want = ceil_div(2 * device_sm_count(), rb);
  if (want > nt) want = nt;
  if (want > max_lists) want = max_lists;
  if (want < 1) want = 1;
  *tps = (int)ceil_div(nt, want);
  *nsplit = (int)ceil_div(nt, *tps);
}

static size_t topk_smem_bytes() {
  return sizeof(float) * (2 * kBK * (kBM + kPad) + kBM * (kBN + 1) + kBM * kMaxK) + sizeof(int64_t) * kBM * kMaxK;
}

}  // namespace pclip

using namespace pclip;

size_t pclip::simt_topk_workspace(int64_t Q, int64_t N, int k, bool list_mode) {
  int nsplit, tps;
  plan_n_split(list_mode ? 64 : Q, N, list_mode ? kMaxLists : kMaxK, &nsplit, &tps);
  if (nsplit == 1 && !list_mode) return 0;
  const size_t entries = list_mode ? (size_t)list_mode_entries(Q, k) : (size_t)nsplit * Q * k;
  return align_up(entries * sizeof(int64_t), 1024) + align_up(entries * sizeof(float), 1024) + 4096;
}

int pclip::simt_topk(const float* q, const float* bank, int64_t Q, int64_t N, int D, int k, int64_t index_base,
                     int64_t* idx, float* score, void* ws, size_t ws_bytes, cudaStream_t st, const int32_t* qlist,
                     const int32_t* qcount) {
  int nsplit, tps;
  // list mode: the number of live rows is only known on the device; plan the bank split for a handful of rows
  plan_n_split(qlist ? 64 : Q, N, qlist ? kMaxLists : kMaxK, &nsplit, &tps);
  const int64_t entries = qlist ? list_mode_entries(Q, k) : (int64_t)nsplit * Q * k;
  int64_t* pidx = idx;
  float* pscore = score;
  if (nsplit > 1 || qlist) {
    Workspace w(ws, ws_bytes);
    PCLIP_REQUIRE(ws, PCLIP_ERR_WORKSPACE, "topk: workspace required");
    pidx = w.take<int64_t>((size_t)entries);
    pscore = w.take<float>((size_t)entries);
    PCLIP_REQUIRE(w.ok(), PCLIP_ERR_WORKSPACE, "topk: workspace too small");
  }
  const size_t smem = topk_smem_bytes();
  PCLIP_CHECK_CUDA(cudaFuncSetAttribute(topk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int64_t row_blocks = ceil_div(Q, kBM);
  dim3 grid((unsigned)(qlist ? std::min<int64_t>(row_blocks, 8) : row_blocks), (unsigned)nsplit);
  topk_kernel<<<grid, 256, smem, st>>>(q, bank, Q, N, D, k, tps, index_base, pidx, pscore, qlist, qcount, entries);
  PCLIP_CHECK_LAUNCH();
  if (nsplit > 1 || qlist) {
    topk_merge_kernel<<<(unsigned)ceil_div(Q, 128), 128, 0, st>>>(pidx, pscore, Q, nsplit, k, idx, score, qlist,
                                                                  qcount, entries, ceil_div(N, kBN));
    PCLIP_CHECK_LAUNCH();
  }
  return PCLIP_OK;
}

extern "C" {

int pclip_scores(const float* q, const float* bank, int64_t Q, int64_t N, int32_t D, int32_t op, const float* neg_tau,
                 float* out, pclip_stream_t stream) {
  PCLIP_TRY(require_device());
  PCLIP_REQUIRE(q && bank && out && Q > 0 && N > 0 && D > 0, PCLIP_ERR_INVALID, "scores: bad argument");
  PCLIP_REQUIRE(op == PCLIP_SCORE_DOT || op == PCLIP_SCORE_EXP, PCLIP_ERR_INVALID, "scores: unknown op %d", op);
  dim3 grid((unsigned)ceil_div(N, kBN), (unsigned)ceil_div(Q, kBM));
  scores_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(q, bank, Q, N, D, op, neg_tau, out);
  PCLIP_CHECK_LAUNCH();
  return PCLIP_OK;
}

size_t pclip_rank_workspace_bytes(int64_t Q) { return align_up((size_t)Q * sizeof(float), 1024); }

int pclip_rank_of_positive(const float* q, const float* bank, const int64_t* pos, int64_t Q, int64_t N, int32_t D,
                           int32_t* rank, void* workspace, size_t workspace_bytes, pclip_stream_t stream) {
  PCLIP_TRY(require_device());
  PCLIP_REQUIRE(q && bank && pos && rank && Q > 0 && N > 0 && D > 0, PCLIP_ERR_INVALID, "rank: bad argument");
  PCLIP_REQUIRE(workspace && (reinterpret_cast<uintptr_t>(workspace) & 1023) == 0 &&
                    workspace_bytes >= pclip_rank_workspace_bytes(Q),
                PCLIP_ERR_WORKSPACE, "rank: workspace missing, misaligned or too small");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  float* pos_score = static_cast<float*>(workspace);   // the positives' scores, [Q]
  pos_score_kernel<<<(unsigned)ceil_div(Q, 128), 128, 0, st>>>(q, bank, pos, Q, D, pos_score, rank);
  PCLIP_CHECK_LAUNCH();
  int nsplit, tps;
  plan_n_split(Q, N, 1 << 20, &nsplit, &tps);
  dim3 grid((unsigned)ceil_div(Q, kBM), (unsigned)nsplit);
  rank_kernel<<<grid, 256, 0, st>>>(q, bank, pos_score, Q, N, D, tps, rank);
  PCLIP_CHECK_LAUNCH();
  return PCLIP_OK;
}

size_t pclip_topk_workspace_bytes(int64_t Q, int64_t N, int32_t D, int32_t k, int32_t compute) {
  (void)D;
  if (compute == PCLIP_COMPUTE_BF16) return tc_topk_workspace(Q, N, k);
  return simt_topk_workspace(Q, N, k, false);
}

int pclip_topk(const float* q, const float* bank, int64_t Q, int64_t N, int32_t D, int32_t k, int32_t compute,
               int64_t index_base, int64_t* idx, float* score, int32_t* n_unsafe, void* workspace,
               size_t workspace_bytes, pclip_stream_t stream) {
  PCLIP_TRY(require_device());
  PCLIP_REQUIRE(q && bank && idx && score && Q > 0 && N > 0 && D > 0, PCLIP_ERR_INVALID, "topk: bad argument");
  PCLIP_REQUIRE(k >= 1 && k <= kMaxK, PCLIP_ERR_UNSUPPORTED, "topk: k must be in [1, %d]", kMaxK);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (compute == PCLIP_COMPUTE_BF16)
    return tc_topk(q, bank, Q, N, D, k, index_base, idx, score, n_unsafe, workspace, workspace_bytes, st);
  PCLIP_REQUIRE(compute == PCLIP_COMPUTE_F32, PCLIP_ERR_UNSUPPORTED, "topk: unknown compute mode %d", compute);
  if (n_unsafe) PCLIP_CHECK_CUDA(cudaMemsetAsync(n_unsafe, 0, sizeof(int32_t), st));
  return simt_topk(q, bank, Q, N, D, k, index_base, idx, score, workspace, workspace_bytes, st);
}

size_t pclip_topk_bank_bytes(int64_t N, int32_t D) {
  (void)D;
  return tc_topk_bank_bytes(N);
}

int pclip_topk_prepare_bank(const float* bank, int64_t N, int32_t D, void* prepared, size_t prepared_bytes,
                            pclip_stream_t stream) {
  PCLIP_TRY(require_device());
  PCLIP_REQUIRE(bank && N > 0, PCLIP_ERR_INVALID, "topk_prepare_bank: bad argument");
  return tc_topk_prepare_bank(bank, N, D, prepared, prepared_bytes, static_cast<cudaStream_t>(stream));
}

int pclip_topk_prepared(const float* q, const float* bank, const void* prepared, int64_t Q, int64_t N, int32_t D,
                        int32_t k, int64_t index_base, int64_t* idx, float* score, int32_t* n_unsafe, void* workspace,
                        size_t workspace_bytes, pclip_stream_t stream) {
  PCLIP_TRY(require_device());
  PCLIP_REQUIRE(q && bank && prepared && idx && score && Q > 0 && N > 0 && D > 0, PCLIP_ERR_INVALID,
                "topk_prepared: bad argument");
  PCLIP_REQUIRE(k >= 1 && k <= kMaxK, PCLIP_ERR_UNSUPPORTED, "topk: k must be in [1, %d]", kMaxK);
  return tc_topk(q, bank, Q, N, D, k, index_base, idx, score, n_unsafe, workspace, workspace_bytes,
                 static_cast<cudaStream_t>(stream), prepared);
}

int pclip_topk_merge(const int64_t* idx_in, const float* score_in, int64_t Q, int32_t n_lists, int32_t k,
                     int64_t* idx, float* score, pclip_stream_t stream) {
  PCLIP_TRY(require_device());
  PCLIP_REQUIRE(idx_in && score_in && idx && score && Q > 0, PCLIP_ERR_INVALID, "topk_merge: bad argument");
  PCLIP_REQUIRE(n_lists >= 1 && n_lists <= kMaxK && k >= 1 && k <= kMaxK, PCLIP_ERR_UNSUPPORTED,
                "topk_merge: n_lists and k must be in [1, %d]", kMaxK);
  topk_merge_kernel<<<(unsigned)ceil_div(Q, 128), 128, 0, static_cast<cudaStream_t>(stream)>>>(
      idx_in, score_in, Q, n_lists, k, idx, score, nullptr, nullptr, 0, 0);
  PCLIP_CHECK_LAUNCH();
  return PCLIP_OK;
}

}  // extern "C"
