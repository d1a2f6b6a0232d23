// Cosine / dot-product top-k retrieval on the tensor cores (PCLIP_COMPUTE_BF16).
//   ref: modules/training_utils.py:219-222 (rank of the positive among the candidates), topk.py:47-61 (top-k accuracy);
//   generalised to "top-k of N bank rows for each of Q queries" (BASELINE.json configs[4]).
//
// Three stages:
//   1. to_bf16_norm_kernel   fp32 rows -> bf16 rows (+ row norms, + max bank norm)          HBM-bound pre-pass
//   2. tc_topk_kernel        each CTA keeps 256 queries resident (two 128-row UMMA A tiles), streams the bank through a
//                            TMA ring, scores 128-column tiles with tcgen05.mma into four TMEM accumulators
//                            (2 query tiles x double buffer) and filters them in the epilogue: thread == query row,
//                            sorted candidate list (KC >= 2k + 4 entries: 16, 24 or 32) held in the thread's registers,
//                            its tail is the filter threshold.
//   3. topk_rescore_kernel   candidates are re-scored with the fp32 FFMA chain of the exact path (bit-identical to
//                            PCLIP_COMPUTE_F32), sorted (score desc, index asc) and cut to k.  A query is PROVEN exact
//                            when its k-th fp32 score exceeds (largest discarded bf16 score + eps_q), eps_q being a
//                            rigorous bound of |bf16 score - fp32 score|; otherwise it is appended to a list that the
//                            exact fp32 kernel (scores.cu, list mode) recomputes.  No host synchronisation anywhere.
#include <cuda_bf16.h>
#include <float.h>

#include "common.cuh"
#include "ptx.cuh"

namespace pclip {

namespace tk {
constexpr int kTile = 128;
constexpr int kDim = 128;
constexpr uint32_t kPanelBytes = 128 * 128;
constexpr uint32_t kTileBytes = 2 * kPanelBytes;
constexpr int kQTiles = 2;              // query tiles per CTA (256 queries): halves the bank traffic per score
constexpr int kRows = kQTiles * kTile;
constexpr int kStages = 5;              // bank ring
constexpr int kMaxCand = 32;            // candidate list capacity per query
constexpr int kThreads = 64 + 32 * 4 * kQTiles;  // TMA warp, MMA warp, 8 epilogue warps
constexpr int kEpiThreadsPerTile = 128;
// smem: A (2 tiles) | B ring | barriers
constexpr uint32_t kSmemBytes = kTileBytes * (kQTiles + kStages) + 256 + 1024;
static_assert(kSmemBytes <= 232448, "shared memory budget");

// K-major tile descriptor, built once per tile and advanced per k-step by a constant (ptx.cuh desc_advance)
__device__ __forceinline__ uint64_t tile_desc_base(uint32_t tile_smem_addr) { return make_smem_desc_sw128(tile_smem_addr, 16, 1024); }
__device__ __forceinline__ uint64_t tile_desc_step(uint64_t base, int k) {
  return desc_advance(base, (uint32_t)(k >> 2) * kPanelBytes + (uint32_t)(k & 3) * 32u);
}
__device__ __forceinline__ float max3(float a, float b, float c) {
  float r;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
  return r;
}
}  // namespace tk

// ---------------------------------------------------------------------------------------------------------
// 1. fp32 [rows, 128] -> bf16 [rows, 128]; norm[r] = ||row r||_2 (optional); *maxnorm = max_r ||row r|| (optional,
//    as an int-ordered float: norms are >= 0).  One warp per row, one float4 per lane.
// ---------------------------------------------------------------------------------------------------------
// Per row: bf16 copy, ||x|| (norm), ||x - bf16(x)|| (err), and for the bank the maxima over all rows of ||bf16(x)||
// and of ||x - bf16(x)|| (max_bits[0], max_bits[1]; non-negative floats order like their bit patterns).  These are the
// MEASURED rounding errors the exactness bound of topk_rescore_kernel is built from.
__global__ void __launch_bounds__(256) to_bf16_norm_kernel(const float* __restrict__ in, int64_t rows,
                                                           __nv_bfloat16* __restrict__ out, float* __restrict__ norm,
                                                           float* __restrict__ err, int* __restrict__ max_bits) {
  const int lane = threadIdx.x & 31;
  const int64_t warp0 = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  float wmax = 0.f, emax = 0.f;
  for (int64_t r = warp0; r < rows; r += nwarps) {
    const float4 v = *reinterpret_cast<const float4*>(in + r * tk::kDim + lane * 4);
    __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
    uint2 pk;
    pk.x = *reinterpret_cast<uint32_t*>(&lo);
    pk.y = *reinterpret_cast<uint32_t*>(&hi);
    *reinterpret_cast<uint2*>(out + r * tk::kDim + lane * 4) = pk;
    const float2 l2 = __bfloat1622float2(lo), h2 = __bfloat1622float2(hi);
    float s = v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
    float sb = l2.x * l2.x + l2.y * l2.y + h2.x * h2.x + h2.y * h2.y;
    const float dx = v.x - l2.x, dy = v.y - l2.y, dz = v.z - h2.x, dw = v.w - h2.y;   // exact in fp32
    float se = dx * dx + dy * dy + dz * dz + dw * dw;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      s += __shfl_xor_sync(0xffffffffu, s, o);
      sb += __shfl_xor_sync(0xffffffffu, sb, o);
      se += __shfl_xor_sync(0xffffffffu, se, o);
    }
    if (lane == 0) {
      if (norm) norm[r] = sqrtf(s);
      if (err) err[r] = sqrtf(se);
    }
    wmax = fmaxf(wmax, sqrtf(sb));
    emax = fmaxf(emax, sqrtf(se));
  }
  if (max_bits && lane == 0) {
    if (wmax > 0.f) atomicMax(max_bits, __float_as_int(wmax));
    if (emax > 0.f) atomicMax(max_bits + 1, __float_as_int(emax));
  }
}

// ---------------------------------------------------------------------------------------------------------
// 2. candidate selection
// ---------------------------------------------------------------------------------------------------------
struct TopkArgs {
  int64_t Q, N;
  int kc;               // candidates kept per (query, bank split)
  int tiles_per_split;  // bank tiles (of 128 rows) per blockIdx.y
  int32_t* cand_idx;    // [nsplit][Q][kc]   bank row (local to this call), -1 = empty slot
  float* cand_score;    // [nsplit][Q][kc]   bf16-path score
};

// r[e] for a runtime e without spilling r to local memory: a 5-level select tree
__device__ __forceinline__ float sel32(const uint32_t (&r)[32], int e) {
  uint32_t a[16], b[8], c[4];
#pragma unroll
  for (int i = 0; i < 16; ++i) a[i] = (e & 1) ? r[2 * i + 1] : r[2 * i];
#pragma unroll
  for (int i = 0; i < 8; ++i) b[i] = (e & 2) ? a[2 * i + 1] : a[2 * i];
#pragma unroll
  for (int i = 0; i < 4; ++i) c[i] = (e & 4) ? b[2 * i + 1] : b[2 * i];
  const uint32_t d0 = (e & 8) ? c[1] : c[0], d1 = (e & 8) ? c[3] : c[2];
  return __uint_as_float((e & 16) ? d1 : d0);
}

template <int KC>
__global__ void __launch_bounds__(tk::kThreads, 1)
tc_topk_kernel(const __grid_constant__ CUtensorMap tq, const __grid_constant__ CUtensorMap tb, const TopkArgs args) {
  using namespace tk;
  const int64_t row0 = (int64_t)blockIdx.x * kRows;
  const int ntiles_all = (int)((args.N + kTile - 1) / kTile);
  const int tile_begin = blockIdx.y * args.tiles_per_split;
  const int tile_end = min(tile_begin + args.tiles_per_split, ntiles_all);
  const int ntiles = max(tile_end - tile_begin, 0);

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* a_tiles = smem;
  uint8_t* b_tiles = a_tiles + kQTiles * kTileBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(b_tiles + kStages * kTileBytes);
  uint64_t* a_full = bars;                       // 1
  uint64_t* b_full = a_full + 1;                 // kStages
  uint64_t* b_empty = b_full + kStages;          // kStages
  uint64_t* t_full = b_empty + kStages;          // [kQTiles][2]
  uint64_t* t_empty = t_full + 2 * kQTiles;      // [kQTiles][2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(t_empty + 2 * kQTiles);

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);   // provably warp-uniform: the MMA warp's loop state lives in uniform registers
  const int lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    mbar_init(a_full, 1);
    for (int i = 0; i < kStages; ++i) {
      mbar_init(&b_full[i], 1);
      mbar_init(&b_empty[i], 1);
    }
    for (int i = 0; i < 2 * kQTiles; ++i) {
      mbar_init(&t_full[i], 1);
      mbar_init(&t_empty[i], kEpiThreadsPerTile);
    }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc<512>(tmem_slot);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      tma_prefetch_desc(&tq);
      tma_prefetch_desc(&tb);
      mbar_arrive_expect_tx(a_full, kQTiles * kTileBytes);
      for (int h = 0; h < kQTiles; ++h) {
        tma_load_2d(a_tiles + h * kTileBytes, &tq, a_full, 0, (int32_t)(row0 + h * kTile));
        tma_load_2d(a_tiles + h * kTileBytes + kPanelBytes, &tq, a_full, 64, (int32_t)(row0 + h * kTile));
      }
      for (int t = 0; t < ntiles; ++t) {
        const int stage = t % kStages;
        const uint32_t phase = (uint32_t)(t / kStages) & 1u;
        mbar_wait(&b_empty[stage], phase ^ 1u);
        mbar_arrive_expect_tx(&b_full[stage], kTileBytes);
        const int32_t col0 = (tile_begin + t) * kTile;
        uint8_t* dst = b_tiles + (size_t)stage * kTileBytes;
        tma_load_2d(dst, &tb, &b_full[stage], 0, col0);
        tma_load_2d(dst + kPanelBytes, &tb, &b_full[stage], 64, col0);
      }
    }
  } else if (warp == 1) {
    {   // the whole warp runs the issue loop; one elected lane issues each tcgen05 instruction
      constexpr uint32_t idesc = make_idesc_bf16(kTile, kTile, 0, 0);
      mbar_wait(a_full, 0);
      for (int t = 0; t < ntiles; ++t) {
        const int stage = t % kStages;
        const uint32_t phase = (uint32_t)(t / kStages) & 1u;
        const int buf = t & 1;
        const uint32_t bphase = (uint32_t)(t >> 1) & 1u;
        mbar_wait(&b_full[stage], phase);
        const uint64_t bd = tile_desc_base(smem_u32(b_tiles + (size_t)stage * kTileBytes));
#pragma unroll
        for (int h = 0; h < kQTiles; ++h) {
          mbar_wait(&t_empty[h * 2 + buf], bphase ^ 1u);
          tc_fence_after_sync();
          const uint64_t ad = tile_desc_base(smem_u32(a_tiles + h * kTileBytes));
          const uint32_t d = tmem_base + (uint32_t)(h * 2 + buf) * kTile;
          if (elect_one()) {
#pragma unroll
            for (int k = 0; k < kDim / 16; ++k)
              umma_bf16_ss(d, tile_desc_step(ad, k), tile_desc_step(bd, k), idesc, k > 0 ? 1u : 0u);
            umma_commit(&t_full[h * 2 + buf]);
            if (h == kQTiles - 1) umma_commit(&b_empty[stage]);
          }
          __syncwarp();
        }
      }
    }
  } else {
    // ---- epilogue: thread == query row; its sorted candidate list (KC scores + bank rows) lives in registers ----
    const int h = (warp - 2) >> 2;   // query tile
    const int q = warp & 3;          // TMEM lane quadrant this warp may read
    const int lrow = h * kTile + q * 32 + lane;
    float cs[KC];                    // descending; -inf = empty slot
    int32_t ci[KC];
#pragma unroll
    for (int j = 0; j < KC; ++j) {
      cs[j] = -INFINITY;
      ci[j] = -1;
    }
    for (int t = 0; t < ntiles; ++t) {
      const int buf = t & 1;
      const uint32_t bphase = (uint32_t)(t >> 1) & 1u;
      const int64_t col0 = (int64_t)(tile_begin + t) * kTile;
      const bool ragged = col0 + kTile > args.N;
      mbar_wait(&t_full[h * 2 + buf], bphase);
      tc_fence_after_sync();
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(h * 2 + buf) * kTile;
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        uint32_t r[32];
        tmem_ld_32x32b_x32(taddr + c * 32, r);
        tmem_ld_wait();
        if (ragged) {
#pragma unroll
          for (int e = 0; e < 32; ++e)
            if (col0 + c * 32 + e >= args.N) r[e] = __float_as_uint(-FLT_MAX);  // TMA zero-filled rows beyond N
        }
        float m = -FLT_MAX;
#pragma unroll
        for (int e = 0; e < 32; e += 2) m = max3(m, __uint_as_float(r[e]), __uint_as_float(r[e + 1]));
        if (m > cs[KC - 1]) {
          uint32_t hits = 0;   // elements above the threshold as of now; re-checked at insert time
#pragma unroll
          for (int e = 0; e < 32; ++e) hits |= (__uint_as_float(r[e]) > cs[KC - 1] ? 1u : 0u) << e;
#pragma unroll 1
          for (; hits; hits &= hits - 1) {   // ascending e: equal scores keep bank order
            const int e = __ffs(hits) - 1;
            const float v = sel32(r, e);
            if (v > cs[KC - 1] && v > -FLT_MAX) {
              // sorted insert (descending) as a branch-free shift chain over the register list; an equal score stays
              // behind earlier (lower-index) entries
              const int32_t id = (int32_t)(col0 + c * 32 + e);
#pragma unroll
              for (int j = KC - 1; j >= 1; --j) {
                const bool shift = cs[j - 1] < v;
                const bool here = !shift && cs[j] < v;
                cs[j] = shift ? cs[j - 1] : (here ? v : cs[j]);
                ci[j] = shift ? ci[j - 1] : (here ? id : ci[j]);
              }
              if (cs[0] < v) {
                cs[0] = v;
                ci[0] = id;
              }
            }
          }
        }
      }
      tc_fence_before_sync();
      mbar_arrive(&t_empty[h * 2 + buf]);
    }
    const int64_t row = row0 + lrow;
    if (row < args.Q) {
      const int64_t o = ((int64_t)blockIdx.y * args.Q + row) * KC;
#pragma unroll
      for (int j = 0; j < KC; ++j) {
        args.cand_idx[o + j] = ci[j];
        args.cand_score[o + j] = ci[j] >= 0 ? cs[j] : -FLT_MAX;
      }
    }
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after_sync();
    __syncwarp();
    tmem_dealloc<512>(tmem_base);
  }
}

// ---------------------------------------------------------------------------------------------------------
// 3. exact re-scoring of the candidates + proof of exactness.  One warp per query.
// ---------------------------------------------------------------------------------------------------------
constexpr int kRescoreKeep = 64;      // candidates that get an fp32 score when a query has more than that many
constexpr int kRescoreWarps = 8;      // at most; fewer when the candidate lists of a warp's query need more shared memory
constexpr int kMaxCandTotal = 4096;   // nsplit * kc: up to 170 bank splits at kc = 24, so that a call with few queries
                                      // (one query block) still spreads the bank over every SM

// shared memory of one warp: its query (128 floats) and the re-scored candidates (score, index) of that query
static size_t rescore_warp_bytes(int M) { return (size_t)tk::kDim * 4 + (size_t)M * 8 + (size_t)kRescoreKeep * 8; }
static int rescore_warps(int M) {
  return (int)std::max<size_t>(1, std::min<size_t>(kRescoreWarps, (size_t)96 * 1024 / rescore_warp_bytes(M)));
}

__global__ void __launch_bounds__(kRescoreWarps * 32)
topk_rescore_kernel(const float* __restrict__ q, const float* __restrict__ bank, const int32_t* __restrict__ cand_idx,
                    const float* __restrict__ cand_score, const float* __restrict__ qnorm,
                    const float* __restrict__ qerr, const int* __restrict__ bmax_bits, int64_t Q, int nsplit, int kc, int k, int64_t index_base,
                    int64_t* __restrict__ idx_out, float* __restrict__ score_out, int32_t* __restrict__ unsafe_list,
                    int32_t* __restrict__ unsafe_count) {
  extern __shared__ __align__(16) uint8_t rescore_smem[];
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + w;
  if (row >= Q) return;
  const int M = nsplit * kc;
  float* qs = reinterpret_cast<float*>(rescore_smem +
                                       (size_t)w * ((size_t)tk::kDim * 4 + (size_t)M * 8 + (size_t)kRescoreKeep * 8));
  float* cs = qs + tk::kDim;
  int32_t* ci = reinterpret_cast<int32_t*>(cs + M);
  *reinterpret_cast<float4*>(&qs[lane * 4]) = *reinterpret_cast<const float4*>(q + row * tk::kDim + lane * 4);
  __syncwarp();
  float tdisc = -FLT_MAX;  // largest bf16 score any discarded bank row can have
  // (A) all candidates of the query with their bf16 scores; a full list's tail bounds what its split discarded
  for (int c = lane; c < M; c += 32) {
    const int s = c / kc, j = c - s * kc;
    const int64_t o = ((int64_t)s * Q + row) * kc + j;
    const int32_t id = cand_idx[o];
    cs[c] = id >= 0 ? cand_score[o] : -FLT_MAX;
    ci[c] = id >= 0 ? id : 0x7fffffff;
    if (id >= 0 && j == kc - 1) tdisc = fmaxf(tdisc, cand_score[o]);
  }
  __syncwarp();
  // (B) many bank splits (few queries per call): keep only the kRescoreKeep best candidates by bf16 score (ties: lower
  //     index) for the fp32 pass -- the best score among the rest joins the discarded bound, so the proof below still
  //     covers every bank row that is not re-scored.  Every split's list is sorted (best first), so this is a
  //     tournament over the list heads: lane l owns the heads of splits l, l + 32, ...
  int Mr = M;
  if (M > kRescoreKeep) {
    float* ks = reinterpret_cast<float*>(ci + M);
    int32_t* ki = reinterpret_cast<int32_t*>(ks + kRescoreKeep);
    constexpr int kHeads = (kMaxCandTotal / 16 + 31) / 32;      // splits per lane at the smallest list length
    int hp[kHeads];
#pragma unroll
    for (int i = 0; i < kHeads; ++i) hp[i] = 0;
    int kept = 0;
    for (; kept < kRescoreKeep; ++kept) {
      float bs = -FLT_MAX;
      int32_t bi = 0x7fffffff;
      int bsplit = -1;
#pragma unroll
      for (int i = 0; i < kHeads; ++i) {
        const int sp = lane + 32 * i;
        if (sp < nsplit && hp[i] < kc) {
          const float sc = cs[sp * kc + hp[i]];
          const int32_t id = ci[sp * kc + hp[i]];
          if (id != 0x7fffffff && (bsplit < 0 || sc > bs || (sc == bs && id < bi))) bs = sc, bi = id, bsplit = sp;
        }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const float os = __shfl_xor_sync(0xffffffffu, bs, o);
        const int32_t oi = __shfl_xor_sync(0xffffffffu, bi, o);
        const int osp = __shfl_xor_sync(0xffffffffu, bsplit, o);
        if (osp >= 0 && (bsplit < 0 || os > bs || (os == bs && oi < bi))) bs = os, bi = oi, bsplit = osp;
      }
      if (bsplit < 0) break;                   // all lists exhausted (warp-uniform)
#pragma unroll
      for (int i = 0; i < kHeads; ++i)
        if (bsplit == lane + 32 * i) ++hp[i];
      if (lane == 0) ks[kept] = bs, ki[kept] = bi;
    }
    // what was not kept: every split's remaining head is its best
#pragma unroll
    for (int i = 0; i < kHeads; ++i) {
      const int sp = lane + 32 * i;
      if (sp < nsplit && hp[i] < kc && ci[sp * kc + hp[i]] != 0x7fffffff) tdisc = fmaxf(tdisc, cs[sp * kc + hp[i]]);
    }
    __syncwarp();
    for (int c = lane; c < kRescoreKeep; c += 32) {
      cs[c] = c < kept ? ks[c] : -FLT_MAX;
      ci[c] = c < kept ? ki[c] : 0x7fffffff;
    }
    __syncwarp();
    Mr = kRescoreKeep;
  }
  // (C) exact fp32 scores of the kept candidates
  for (int c = lane; c < Mr; c += 32) {
    const int32_t id = ci[c];
    float acc = -FLT_MAX;
    if (id != 0x7fffffff) {
      const float* b = bank + (int64_t)id * tk::kDim;
      acc = 0.f;
#pragma unroll 4
      for (int d = 0; d < tk::kDim; d += 4) {  // the exact path's chain: ascending d, one FFMA per element
        const float4 bv = *reinterpret_cast<const float4*>(b + d);
        acc = fmaf(qs[d], bv.x, acc);
        acc = fmaf(qs[d + 1], bv.y, acc);
        acc = fmaf(qs[d + 2], bv.z, acc);
        acc = fmaf(qs[d + 3], bv.w, acc);
      }
    }
    cs[c] = acc;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) tdisc = fmaxf(tdisc, __shfl_xor_sync(0xffffffffu, tdisc, o));
  __syncwarp();
  float kth = FLT_MAX;
  bool have_k = true;
  for (int p = 0; p < k; ++p) {
    float bs = -FLT_MAX;
    int32_t bi = 0x7fffffff;
    int bc = -1;
    for (int c = lane; c < Mr; c += 32) {
      const float s = cs[c];
      const int32_t id = ci[c];
      if (id != 0x7fffffff && (bc < 0 || s > bs || (s == bs && id < bi))) {
        bs = s;
        bi = id;
        bc = c;
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float os = __shfl_xor_sync(0xffffffffu, bs, o);
      const int32_t oi = __shfl_xor_sync(0xffffffffu, bi, o);
      const int oc = __shfl_xor_sync(0xffffffffu, bc, o);
      if (oc >= 0 && (bc < 0 || os > bs || (os == bs && oi < bi))) {
        bs = os;
        bi = oi;
        bc = oc;
      }
    }
    if (bc >= 0 && (bc & 31) == lane) ci[bc] = 0x7fffffff;  // taken
    __syncwarp();
    if (lane == 0) {
      idx_out[row * k + p] = bc >= 0 ? (int64_t)bi + index_base : -1;
      score_out[row * k + p] = bc >= 0 ? bs : -FLT_MAX;
    }
    if (bc >= 0) kth = bs; else have_k = false;
  }
  if (lane == 0 && tdisc > -FLT_MAX) {
    // q.b - q^.b^ = q.(b - b^) + (q - q^).b^  (^ = rounded to bf16), so by Cauchy-Schwarz
    //   |bf16-path score - fp32-chain score| <= ||q|| max||b - b^|| + ||q - q^|| max||b^|| + D 2^-23 ||q|| max||b^||
    // with the rounding-error norms MEASURED when the operands were converted (to_bf16_norm_kernel) and the last term
    // covering the fp32 accumulation on both sides.  The a-priori form of the same bound is 2^-7 (1 + 2^-9) ||q|| ||b||
    // (each operand moves by at most 2^-8 relative; round 1 used half of that); the measured norms are
    // ~2.4x smaller on random data, which keeps the exact-fallback list short.  1.001 absorbs the fp32 rounding of the
    // norms themselves.
    const float bhat = __int_as_float(bmax_bits[0]), berr = __int_as_float(bmax_bits[1]);
    const float eps = 1.001f * (qnorm[row] * berr + qerr[row] * bhat) + 128.0f * 1.1920929e-7f * 1.01f * qnorm[row] * bhat;
    if (!have_k || !(kth > tdisc + eps)) unsafe_list[atomicAdd(unsafe_count, 1)] = (int32_t)row;
  }
}

// ---------------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------------
// candidates kept per (query, split): >= 2k + 4, rounded up to a list length the kernel is instantiated for
static int cand_per_query(int k) { return k <= 6 ? 16 : (k <= 10 ? 24 : tk::kMaxCand); }

// Bank splits: fill whole waves of the SMs (one CTA per SM).  Cost model as for the CLIP kernels: waves x (tiles per
// split + overhead), overhead = a CTA's start-up phase (lists still filling, most tiles insert) in units of one bank
// tile.  512 query blocks (131 072 queries): 2 splits = 6.9 waves of 16 384 tiles instead of 3.5 -> 4 waves of 32 768.
static void plan_topk_split(int64_t Q, int64_t N, int kc, int* nsplit, int* tps) {
  const int64_t rb = ceil_div(Q, tk::kRows), nt = ceil_div(N, tk::kTile);
  const int64_t sms = std::max(1, device_sm_count()), overhead = 128;
  const int64_t max_ns = std::max<int64_t>(1, std::min<int64_t>(kMaxCandTotal / kc, nt));
  int64_t best_ns = 1, best_cost = -1;
  for (int64_t ns = 1; ns <= max_ns; ++ns) {
    const int64_t t = ceil_div(nt, ns);
    const int64_t cost = ceil_div(rb * ceil_div(nt, t), sms) * (t + overhead);
    if (best_cost < 0 || cost < best_cost) {
      best_cost = cost;
      best_ns = ns;
    }
  }
  *tps = (int)ceil_div(nt, best_ns);
  *nsplit = (int)ceil_div(nt, *tps);
}

struct TopkWs {
  __nv_bfloat16 *qb, *bb;
  float *qnorm, *qerr;
  int32_t* scalars;  // [0] bits of max ||bf16(b)||, [1] bits of max ||b - bf16(b)||, [2] unsafe count
  int32_t* unsafe;
  int32_t* cand_idx;
  float* cand_score;
  void* fix;
  size_t fix_bytes;
  bool ok;
};

static TopkWs carve(void* ws, size_t ws_bytes, int64_t Q, int64_t N, int k, int kc, int nsplit) {
  Workspace w(ws, ws_bytes);
  TopkWs t;
  t.qb = w.take<__nv_bfloat16>((size_t)(Q + tk::kRows) * tk::kDim);
  t.bb = w.take<__nv_bfloat16>((size_t)(N + tk::kTile) * tk::kDim);
  t.qnorm = w.take<float>(Q);
  t.qerr = w.take<float>(Q);
  t.scalars = w.take<int32_t>(4);
  t.unsafe = w.take<int32_t>(Q);
  t.cand_idx = w.take<int32_t>((size_t)nsplit * Q * kc);
  t.cand_score = w.take<float>((size_t)nsplit * Q * kc);
  t.fix_bytes = simt_topk_workspace(Q, N, k, true);
  t.fix = w.take<char>(t.fix_bytes);
  t.ok = ws != nullptr && w.ok();
  return t;
}

size_t tc_topk_workspace(int64_t Q, int64_t N, int k) {
  const int kc = cand_per_query(k);
  int nsplit, tps;
  plan_topk_split(Q, N, kc, &nsplit, &tps);
  Workspace w(nullptr, 0);
  w.take<__nv_bfloat16>((size_t)(Q + tk::kRows) * tk::kDim);
  w.take<__nv_bfloat16>((size_t)(N + tk::kTile) * tk::kDim);
  w.take<float>(Q);
  w.take<float>(Q);
  w.take<int32_t>(4);
  w.take<int32_t>(Q);
  w.take<int32_t>((size_t)nsplit * Q * kc);
  w.take<float>((size_t)nsplit * Q * kc);
  w.take<char>(simt_topk_workspace(Q, N, k, true));
  return align_up(w.off, 1024) + 1024;
}

// Prepared bank image: bf16 rows [(N + kTile) x 128] followed (1024-byte aligned) by the bits of max ||bf16(b)|| and of
// max ||b - bf16(b)||.
size_t tc_topk_bank_bytes(int64_t N) {
  return align_up((size_t)(N + tk::kTile) * tk::kDim * sizeof(__nv_bfloat16), 1024) + 1024;
}

int tc_topk_prepare_bank(const float* bank, int64_t N, int D, void* prepared, size_t prepared_bytes, cudaStream_t st) {
  PCLIP_REQUIRE(D == tk::kDim, PCLIP_ERR_UNSUPPORTED, "topk: bf16 tensor-core path needs D == %d (got %d)", tk::kDim, D);
  PCLIP_REQUIRE(prepared && (reinterpret_cast<uintptr_t>(prepared) & 1023) == 0 && prepared_bytes >= tc_topk_bank_bytes(N),
                PCLIP_ERR_WORKSPACE, "topk_prepare_bank: buffer missing, misaligned or smaller than %zu bytes",
                tc_topk_bank_bytes(N));
  PCLIP_REQUIRE((reinterpret_cast<uintptr_t>(bank) & 15) == 0, PCLIP_ERR_INVALID, "topk: bank must be 16-byte aligned");
  int32_t* maxbits = reinterpret_cast<int32_t*>(static_cast<char*>(prepared) + tc_topk_bank_bytes(N) - 1024);
  PCLIP_CHECK_CUDA(cudaMemsetAsync(maxbits, 0, 2 * sizeof(int32_t), st));
  to_bf16_norm_kernel<<<device_sm_count() * 8, 256, 0, st>>>(bank, N, static_cast<__nv_bfloat16*>(prepared), nullptr,
                                                             nullptr, maxbits);
  PCLIP_CHECK_LAUNCH();
  return PCLIP_OK;
}

int tc_topk(const float* q, const float* bank, int64_t Q, int64_t N, int D, int k, int64_t index_base, int64_t* idx,
            float* score, int32_t* n_unsafe, void* ws, size_t ws_bytes, cudaStream_t st, const void* prepared) {
  PCLIP_REQUIRE(D == tk::kDim, PCLIP_ERR_UNSUPPORTED, "topk: bf16 tensor-core path needs D == %d (got %d)", tk::kDim, D);
  PCLIP_REQUIRE(N < (int64_t)0x7fffff00, PCLIP_ERR_UNSUPPORTED, "topk: N must fit 31 bits per call (shard the bank)");
  PCLIP_REQUIRE((reinterpret_cast<uintptr_t>(q) & 15) == 0 && (reinterpret_cast<uintptr_t>(bank) & 15) == 0,
                PCLIP_ERR_INVALID, "topk: q and bank must be 16-byte aligned");
  const int kc = cand_per_query(k);
  PCLIP_REQUIRE(k <= kc, PCLIP_ERR_UNSUPPORTED, "topk: bf16 path supports k <= %d", tk::kMaxCand);
  int nsplit, tps;
  plan_topk_split(Q, N, kc, &nsplit, &tps);
  PCLIP_REQUIRE(ws && (reinterpret_cast<uintptr_t>(ws) & 1023) == 0, PCLIP_ERR_WORKSPACE,
                "topk: workspace must be 1024-byte aligned");
  TopkWs t = carve(ws, ws_bytes, Q, N, k, kc, nsplit);
  PCLIP_REQUIRE(t.ok, PCLIP_ERR_WORKSPACE, "topk: workspace too small (%zu < %zu)", ws_bytes,
                tc_topk_workspace(Q, N, k));
  PCLIP_CHECK_CUDA(cudaMemsetAsync(t.scalars, 0, 4 * sizeof(int32_t), st));
  const int cblocks = device_sm_count() * 8;
  to_bf16_norm_kernel<<<cblocks, 256, 0, st>>>(q, Q, t.qb, t.qnorm, t.qerr, nullptr);
  PCLIP_CHECK_LAUNCH();
  const __nv_bfloat16* bb = t.bb;
  if (prepared) {   // persistent bf16 bank (pclip_topk_prepare_bank): no 3 N D-byte conversion pass per call
    bb = static_cast<const __nv_bfloat16*>(prepared);
    PCLIP_CHECK_CUDA(cudaMemcpyAsync(t.scalars, static_cast<const char*>(prepared) + tc_topk_bank_bytes(N) - 1024,
                                     2 * sizeof(int32_t), cudaMemcpyDeviceToDevice, st));
  } else {
    to_bf16_norm_kernel<<<cblocks, 256, 0, st>>>(bank, N, t.bb, nullptr, nullptr, t.scalars);
    PCLIP_CHECK_LAUNCH();
  }
  CUtensorMap mq, mb;
  PCLIP_TRY(make_tmap_bf16_2d(&mq, t.qb, (uint64_t)Q, tk::kDim, tk::kDim, tk::kTile, 64));
  PCLIP_TRY(make_tmap_bf16_2d(&mb, bb, (uint64_t)N, tk::kDim, tk::kDim, tk::kTile, 64));
  TopkArgs a{Q, N, kc, tps, t.cand_idx, t.cand_score};
  auto kern = kc == 16 ? tc_topk_kernel<16> : (kc == 24 ? tc_topk_kernel<24> : tc_topk_kernel<32>);
  PCLIP_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tk::kSmemBytes));
  dim3 grid((unsigned)ceil_div(Q, tk::kRows), (unsigned)nsplit);
  kern<<<grid, tk::kThreads, tk::kSmemBytes, st>>>(mq, mb, a);
  PCLIP_CHECK_LAUNCH();
  const int rw = rescore_warps(nsplit * kc);
  const size_t rsmem = rw * rescore_warp_bytes(nsplit * kc);
  PCLIP_CHECK_CUDA(cudaFuncSetAttribute(topk_rescore_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rsmem));
  topk_rescore_kernel<<<(unsigned)ceil_div(Q, rw), rw * 32, rsmem, st>>>(
      q, bank, t.cand_idx, t.cand_score, t.qnorm, t.qerr, t.scalars, Q, nsplit, kc, k, index_base, idx, score,
      t.unsafe, t.scalars + 2);
  PCLIP_CHECK_LAUNCH();
  // queries whose exactness could not be proven are recomputed by the exact fp32 kernel (device-side count)
  PCLIP_TRY(simt_topk(q, bank, Q, N, D, k, index_base, idx, score, t.fix, t.fix_bytes, st, t.unsafe, t.scalars + 2));
  if (n_unsafe)
    PCLIP_CHECK_CUDA(cudaMemcpyAsync(n_unsafe, t.scalars + 2, sizeof(int32_t), cudaMemcpyDeviceToDevice, st));
  return PCLIP_OK;
}

}  // namespace pclip
