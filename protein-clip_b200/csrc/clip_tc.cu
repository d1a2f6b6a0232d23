// BF16 tensor-core path of the CLIP InfoNCE contractions (PCLIP_COMPUTE_BF16): tcgen05.mma with TMEM
// accumulators, operands staged by TMA into 128-byte-swizzled shared memory, warp-specialised
// (warp 0 = TMA producer, warp 1 = MMA issuer + TMEM owner, warps 2..5 = epilogue, one per TMEM lane quadrant).
//
// Score tiles never reach HBM.  ref: modules/training_utils.py:85-89 (S = x @ y.T; exp; row / column sums) and
// :260-264 (the same under autocast).  D is fixed at 128 (ref: main.py:27 embedding_dim).
//
//   tc_rowsumexp_kernel : rowsum_i = sum_j exp(a_i . b_j), diag_i = a_i . b_(off+i)
//   (two-sided forward: the production path below 2048 columns, where its fp32-exact column sums matter; the
//   single-pass forward and the gradient kernel live in clip_tc2.cu)
//
// All shared-memory operand tiles are K-major [128 rows x 64 bf16] panels (128 B per row, SWIZZLE_128B).
#include <cuda_bf16.h>

#include "common.cuh"
#include "ptx.cuh"

namespace pclip {

constexpr int kTile = 128;                    // rows per CTA (UMMA M) and columns per tile (UMMA N)
constexpr int kDim = 128;                     // embedding dim (UMMA K total)
constexpr uint32_t kPanelBytes = 128 * 128;   // [128 rows x 64 bf16]
constexpr uint32_t kTileBytes = 2 * kPanelBytes;
constexpr int kThreads = 192;
constexpr int kEpiThreads = 128;
static int plan_sms() { return std::max(1, device_sm_count()); }
constexpr float kLog2e = 1.4426950408889634f;

// ---- UMMA descriptors for a K-major SW128 panel pair ------------------------------------------------------------
// k-step `k` (16 bf16 = 32 B) of a [128 x 128] bf16 tile stored as two [128 x 64] panels.
__device__ __forceinline__ uint64_t tile_desc(uint32_t tile_smem_addr, int k) {
  const uint32_t addr = tile_smem_addr + (uint32_t)(k >> 2) * kPanelBytes + (uint32_t)(k & 3) * 32u;
  return make_smem_desc_sw128(addr, 16, 1024);
}

__device__ __forceinline__ void issue_tile_mma(uint32_t d_tmem, uint32_t a_smem, uint32_t b_smem, uint32_t idesc,
                                               bool accumulate_first) {
#pragma unroll
  for (int k = 0; k < kDim / 16; ++k)
    umma_bf16_ss(d_tmem, tile_desc(a_smem, k), tile_desc(b_smem, k), idesc, (k > 0 || accumulate_first) ? 1u : 0u);
}

// =================================================================================================================
// rowsumexp
// =================================================================================================================
struct TcFwdSide {
  int64_t Ma, Nb, diag_offset;
  float* out;      // [nsplit][Ma] partial sums (== rowsum when nsplit == 1)
  float* diag;     // [Ma] or NULL
};
struct TcFwdArgs {
  TcFwdSide s[2];
  int tiles_per_split;
};

constexpr int kFwdStages = 4;
constexpr uint32_t kFwdSmemBytes = kTileBytes * (1 + kFwdStages) + 256 + 1024;

__global__ void __launch_bounds__(kThreads, 1)
tc_rowsumexp_kernel(const __grid_constant__ CUtensorMap ta0, const __grid_constant__ CUtensorMap tb0,
                    const __grid_constant__ CUtensorMap ta1, const __grid_constant__ CUtensorMap tb1,
                    const TcFwdArgs args) {
  const TcFwdSide& sd = args.s[blockIdx.z];
  const CUtensorMap* ta = blockIdx.z == 0 ? &ta0 : &ta1;
  const CUtensorMap* tb = blockIdx.z == 0 ? &tb0 : &tb1;
  const int64_t row0 = (int64_t)blockIdx.x * kTile;
  if (row0 >= sd.Ma) return;
  const int ntiles_all = (int)((sd.Nb + kTile - 1) / kTile);
  const int tile_begin = blockIdx.y * args.tiles_per_split;
  const int tile_end = min(tile_begin + args.tiles_per_split, ntiles_all);
  const int ntiles = max(tile_end - tile_begin, 0);

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* a_tile = smem;
  uint8_t* b_tiles = smem + kTileBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kTileBytes * (1 + kFwdStages));
  uint64_t* a_full = bars;
  uint64_t* b_full = bars + 1;
  uint64_t* b_empty = b_full + kFwdStages;
  uint64_t* t_full = b_empty + kFwdStages;
  uint64_t* t_empty = t_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(t_empty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    mbar_init(a_full, 1);
    for (int i = 0; i < kFwdStages; ++i) {
      mbar_init(&b_full[i], 1);
      mbar_init(&b_empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&t_full[i], 1);
      mbar_init(&t_empty[i], kEpiThreads);
    }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc<256>(tmem_slot);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      tma_prefetch_desc(ta);
      tma_prefetch_desc(tb);
      mbar_arrive_expect_tx(a_full, kTileBytes);
      tma_load_2d(a_tile, ta, a_full, 0, (int32_t)row0);
      tma_load_2d(a_tile + kPanelBytes, ta, a_full, 64, (int32_t)row0);
      for (int t = 0; t < ntiles; ++t) {
        const int stage = t % kFwdStages;
        const uint32_t phase = (uint32_t)(t / kFwdStages) & 1u;
        mbar_wait(&b_empty[stage], phase ^ 1u);
        mbar_arrive_expect_tx(&b_full[stage], kTileBytes);
        const int32_t col0 = (tile_begin + t) * kTile;
        uint8_t* dst = b_tiles + (size_t)stage * kTileBytes;
        tma_load_2d(dst, tb, &b_full[stage], 0, col0);
        tma_load_2d(dst + kPanelBytes, tb, &b_full[stage], 64, col0);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(kTile, kTile, 0, 0);
      const uint32_t a_addr = smem_u32(a_tile);
      mbar_wait(a_full, 0);
      for (int t = 0; t < ntiles; ++t) {
        const int stage = t % kFwdStages;
        const uint32_t phase = (uint32_t)(t / kFwdStages) & 1u;
        const int buf = t & 1;
        const uint32_t bphase = (uint32_t)(t >> 1) & 1u;
        mbar_wait(&t_empty[buf], bphase ^ 1u);
        mbar_wait(&b_full[stage], phase);
        tc_fence_after_sync();
        issue_tile_mma(tmem_base + (uint32_t)buf * kTile, a_addr, smem_u32(b_tiles + (size_t)stage * kTileBytes),
                       idesc, false);
        umma_commit(&b_empty[stage]);
        umma_commit(&t_full[buf]);
      }
    }
  } else {
    const int q = warp & 3;  // TMEM lane quadrant this warp may read
    const int64_t row = row0 + q * 32 + lane;
    const int64_t drow = sd.diag_offset + row;
    const int64_t dlo = sd.diag_offset + row0, dhi = dlo + kTile;  // diagonal columns of this row block
    float rsum = 0.f, dval = 0.f;
    bool dfound = false;
    for (int t = 0; t < ntiles; ++t) {
      const int buf = t & 1;
      const uint32_t bphase = (uint32_t)(t >> 1) & 1u;
      const int64_t col0 = (int64_t)(tile_begin + t) * kTile;
      const bool special = (col0 + kTile > sd.Nb) || (col0 < dhi && col0 + kTile > dlo);
      mbar_wait(&t_full[buf], bphase);
      tc_fence_after_sync();
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)buf * kTile;
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        uint32_t r[32];
        tmem_ld_32x32b_x32(taddr + c * 32, r);
        tmem_ld_wait();
        if (!special) {
          float s0 = 0.f, s1 = 0.f;
#pragma unroll
          for (int e = 0; e < 32; e += 2) {
            s0 += ex2_approx(__uint_as_float(r[e]) * kLog2e);
            s1 += ex2_approx(__uint_as_float(r[e + 1]) * kLog2e);
          }
          rsum += s0 + s1;
        } else {
#pragma unroll
          for (int e = 0; e < 32; ++e) {
            const int64_t col = col0 + c * 32 + e;
            const float v = __uint_as_float(r[e]);
            if (col < sd.Nb) rsum += ex2_approx(v * kLog2e);
            if (col == drow) {
              dval = v;
              dfound = true;
            }
          }
        }
      }
      tc_fence_before_sync();
      mbar_arrive(&t_empty[buf]);
    }
    if (row < sd.Ma) {
      sd.out[(int64_t)blockIdx.y * sd.Ma + row] = rsum;
      if (dfound && sd.diag) sd.diag[row] = dval;
    }
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after_sync();
    __syncwarp();
    tmem_dealloc<256>(tmem_base);
  }
}

// out[i] = sum_s partial[s][i]
__global__ void sum_partials_kernel(const float* __restrict__ partial, int nsplit, int64_t M, float* __restrict__ out) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= M) return;
  float s = 0.f;
  for (int k = 0; k < nsplit; ++k) s += partial[(int64_t)k * M + i];
  out[i] = s;
}

static void plan_split(int64_t Ma, int64_t Nb, int n_sides, int max_split, int* nsplit, int* tps) {
  const int64_t rb = ceil_div(Ma, kTile), nt = ceil_div(Nb, kTile);
  int64_t want = ceil_div(4 * plan_sms(), rb * n_sides);
  if (want > nt) want = nt;
  if (want > max_split) want = max_split;
  if (want < 1 || nt <= 2) want = 1;  // tiny problems: one launch, no partial-sum pass
  *tps = (int)ceil_div(nt, want);
  *nsplit = (int)ceil_div(nt, *tps);
}

size_t tc_rowsumexp_workspace(int64_t Ma, int64_t Nb) {
  int ns, tps;
  plan_split(Ma, Nb, 1, 64, &ns, &tps);
  return 2 * align_up((size_t)ns * Ma * sizeof(float), 1024) + 2048;
}

int tc_rowsumexp(const pclip_clip_fwd_side_t* sides, int n_sides, int D, void* ws, size_t ws_bytes, cudaStream_t st) {
  PCLIP_REQUIRE(D == kDim, PCLIP_ERR_UNSUPPORTED, "bf16 tensor-core path needs D == %d (got %d)", kDim, D);
  for (int i = 0; i < n_sides; ++i) {
    PCLIP_REQUIRE(sides[i].a && sides[i].b && sides[i].rowsum && sides[i].Ma > 0 && sides[i].Nb > 0,
                  PCLIP_ERR_INVALID, "rowsumexp: bad side %d", i);
    PCLIP_REQUIRE(sides[i].Ma == sides[0].Ma && sides[i].Nb == sides[0].Nb, PCLIP_ERR_UNSUPPORTED,
                  "rowsumexp: both sides must have the same shape");
  }
  const int64_t Ma = sides[0].Ma, Nb = sides[0].Nb;
  int nsplit, tps;
  plan_split(Ma, Nb, n_sides, 64, &nsplit, &tps);
  Workspace w(ws, ws_bytes);
  TcFwdArgs args;
  CUtensorMap maps[4];
  float* partial[2] = {nullptr, nullptr};
  for (int i = 0; i < 2; ++i) {
    const pclip_clip_fwd_side_t& s = sides[i < n_sides ? i : 0];
    PCLIP_TRY(make_tmap_bf16_2d(&maps[2 * i], s.a, (uint64_t)s.Ma, kDim, kDim, kTile, 64));
    PCLIP_TRY(make_tmap_bf16_2d(&maps[2 * i + 1], s.b, (uint64_t)s.Nb, kDim, kDim, kTile, 64));
    if (nsplit > 1 && i < n_sides) {
      PCLIP_REQUIRE(ws, PCLIP_ERR_WORKSPACE, "rowsumexp: workspace required");
      partial[i] = w.take<float>((size_t)nsplit * s.Ma);
    }
    args.s[i] = {s.Ma, s.Nb, s.diag_offset, nsplit > 1 ? partial[i] : s.rowsum, s.diag};
  }
  PCLIP_REQUIRE(nsplit == 1 || w.ok(), PCLIP_ERR_WORKSPACE, "rowsumexp: workspace too small");
  args.tiles_per_split = tps;
  PCLIP_CHECK_CUDA(cudaFuncSetAttribute(tc_rowsumexp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        (int)kFwdSmemBytes));
  dim3 grid((unsigned)ceil_div(Ma, kTile), (unsigned)nsplit, (unsigned)n_sides);
  tc_rowsumexp_kernel<<<grid, kThreads, kFwdSmemBytes, st>>>(maps[0], maps[1], maps[2], maps[3], args);
  PCLIP_CHECK_LAUNCH();
  if (nsplit > 1) {
    for (int i = 0; i < n_sides; ++i) {
      sum_partials_kernel<<<(unsigned)ceil_div(Ma, 256), 256, 0, st>>>(partial[i], nsplit, Ma, sides[i].rowsum);
      PCLIP_CHECK_LAUNCH();
    }
  }
  return PCLIP_OK;
}

// da[i, :] = s * (sum_k partial[k][i, :] - 2 b[pos(i), :]) for the column-split partials of the gradient kernel
// (clip_tc2.cu); both sides of a step in one launch (blockIdx.y), partials added in split order.
__global__ void grad_reduce_kernel(GradReduceSide s0, GradReduceSide s1, int nsplit, int64_t Ma, float scale,
                                   const float* __restrict__ gscale) {
  const GradReduceSide& sd = blockIdx.y == 0 ? s0 : s1;
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;  // float4 index
  if (i >= Ma * (kDim / 4)) return;
  const int64_t row = i / (kDim / 4);
  const int c4 = (int)(i % (kDim / 4)) * 4;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
  for (int k = 0; k < nsplit; ++k) {
    const float4 v = __ldcs(reinterpret_cast<const float4*>(sd.partial + ((int64_t)k * Ma + row) * kDim + c4));
    acc.x += v.x;
    acc.y += v.y;
    acc.z += v.z;
    acc.w += v.w;
  }
  const __nv_bfloat16* bp = static_cast<const __nv_bfloat16*>(sd.b) + (sd.diag_offset + row) * kDim + c4;
  const float s = scale * (gscale ? gscale[0] : 1.0f);
  float4 o;
  o.x = s * (acc.x - 2.0f * __bfloat162float(bp[0]));
  o.y = s * (acc.y - 2.0f * __bfloat162float(bp[1]));
  o.z = s * (acc.z - 2.0f * __bfloat162float(bp[2]));
  o.w = s * (acc.w - 2.0f * __bfloat162float(bp[3]));
  *reinterpret_cast<float4*>(sd.da + row * kDim + c4) = o;
}

}  // namespace pclip
