// FILIP token-wise late interaction on the tensor cores (PCLIP_COMPUTE_BF16, D == 128).
//   ref: modules/models.py:292-356 (_filip_similarity_score: einsum -> mask -> amax over s / over t -> masked means)
// The reference materialises S[1,1,Ba,Bb,TA,TB] (64 GiB at B = 256, T = 512) three times over; here a score tile lives
// only in TMEM.
//
//   filip_prep_kernel      fp32 tokens -> bf16 rows [B,T,128] (forward operands; the backward gathers the same rows),
//                          pads zeroed
//   tc_filip_max_kernel    one "side" (see filip.cu): CTA = 256 tokens of one sample (two UMMA row tiles sharing the
//                          streamed column tiles), TMA ring -> tcgen05.mma -> 4 TMEM accumulators; epilogue thread ==
//                          token row keeps the running max / argmax over the column sample's tokens in registers,
//                          applies the length mask, and emits int16 argmax + per-warp partial sums of the masked mean
//   backward               filip_bwd.cu (gather / scatter over the stored argmax indices)
#include <cuda_bf16.h>
#include <float.h>

#include "common.cuh"
#include "ptx.cuh"

namespace pclip {

namespace fl {
constexpr int kTile = 128;
constexpr int kDim = 128;
constexpr uint32_t kPanelBytes = 128 * 128;
constexpr uint32_t kTileBytes = 2 * kPanelBytes;
constexpr int kH = 2;                       // row tiles per CTA
constexpr int kRows = kH * kTile;
constexpr int kThreads = 64 + 32 * 4 * kH;  // TMA warp, MMA warp, 4 warps per row tile

// One 128 x 128 x 128 tile product.  Called by the whole (converged) MMA warp: the descriptors are built once per tile
// and advanced by constants per k-step, and one elected lane issues, so the operands stay in uniform registers (a
// descriptor rebuilt per MMA behind `if (lane == 0)` paced these MMAs at 85-113 clk instead of 64, tools/probe_issue.cu).
__device__ __forceinline__ void issue_tile(uint32_t d, uint32_t a_addr, uint32_t b_addr, uint32_t idesc, bool acc) {
  const uint64_t ad = make_smem_desc_sw128(a_addr, 16, 1024), bd = make_smem_desc_sw128(b_addr, 16, 1024);
  if (elect_one()) {
#pragma unroll
    for (int k = 0; k < kDim / 16; ++k) {
      const uint32_t off = (uint32_t)(k >> 2) * kPanelBytes + (uint32_t)(k & 3) * 32u;
      umma_bf16_ss(d, desc_advance(ad, off), desc_advance(bd, off), idesc, (k > 0 || acc) ? 1u : 0u);
    }
  }
  __syncwarp();
}
}  // namespace fl

// ---------------------------------------------------------------------------------------------------------
// prep: one warp per token row (grid-stride), lane = 4 dims
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
filip_prep_kernel(const float* __restrict__ h, const int32_t* __restrict__ len, int64_t B, int T,
                  __nv_bfloat16* __restrict__ rows /* [B,T,128] */) {
  const int lane = threadIdx.x & 31;
  const int64_t nrow = B * T;
  for (int64_t r = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); r < nrow; r += (int64_t)gridDim.x * 8) {
    const int b = (int)(r / T), t = (int)(r - (int64_t)b * T);
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (t < len[b]) v = __ldg(reinterpret_cast<const float4*>(h + r * fl::kDim) + lane);
    __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
    uint2 pk;
    pk.x = *reinterpret_cast<uint32_t*>(&lo);
    pk.y = *reinterpret_cast<uint32_t*>(&hi);
    *(reinterpret_cast<uint2*>(rows + r * fl::kDim) + lane) = pk;
  }
}

// ---------------------------------------------------------------------------------------------------------
// forward side: grid (ceil(Tx/256), Bx, ysplit)
// ---------------------------------------------------------------------------------------------------------
struct FilipFwdArgs {
  int Bx, By, Tx, Ty;
  const int32_t* lenX;
  const int32_t* lenY;
  int y_per_split;
  float* part;   // [Bx][By][np], np = ceil(Tx/256) * 8 (one slot per epilogue warp)
  int np;
  int16_t* arg;  // [Bx][By][Tx]
};

namespace fl {
constexpr int kFwdStages = 4;
constexpr uint32_t kFwdSmem = kTileBytes * (kH + kFwdStages) + 256 + 1024;
}  // namespace fl

namespace fl {
constexpr uint32_t kNegKey = 0xff7fffe0u;   // key of -FLT_MAX

template <int E>
__device__ __forceinline__ void make_keys(const uint32_t (&r)[32], const uint32_t keep, float (&k)[32]) {
  if constexpr (E < 32) {
    uint32_t key;
    asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(key) : "r"(r[E]), "r"(keep), "n"(31 - E));   // (r & keep) | (31 - E)
    k[E] = __uint_as_float(key);
    make_keys<E + 1>(r, keep, k);
  }
}

__device__ __forceinline__ float max3(float a, float b, float c) { return fmaxf(fmaxf(a, b), c); }   // one FMNMX3

// Maximum and arg-maximum of 32 scores of one token row in 1.5 instructions per score: every score becomes a KEY (its
// 5 low mantissa bits replaced by 31 - column, one LOP3), the keys go through a 3-input max tree (16 FMNMX3), and
// the winner is compared with the running key on the score bits only, so that among scores equal to 2^-18 relative
// the lowest column of the chunk wins and an earlier chunk beats a later one (non-negative maxima; the reference's
// amax backward splits exact ties).  The round-1 epilogue (compare, select, select per score, one
// 64-deep dependent chain per chunk) paced the kernel at 1160 clk per tile pair against 1024 clk of MMA: 8.0 ms per
// forward at cfg 3, 6.8 ms with this one.  Making half of the keys on the FMA pipe (IMAD.HI + IMAD) was slower (7.5 ms).
__device__ __forceinline__ void chunk_max(const uint32_t (&r)[32], const int col0, uint32_t& mk, int& base) {
  float k[32];
  uint32_t keep;   // 0xffffffe0 in a register the optimiser cannot see through: (r & keep) | imm is then ONE LOP3
  asm("mov.u32 %0, 0xffffffe0;" : "=r"(keep));
  make_keys<0>(r, keep, k);
  float a[11];
#pragma unroll
  for (int j = 0; j < 10; ++j) a[j] = max3(k[3 * j], k[3 * j + 1], k[3 * j + 2]);
  a[10] = fmaxf(k[30], k[31]);
  const float b0 = max3(a[0], a[1], a[2]), b1 = max3(a[3], a[4], a[5]), b2 = max3(a[6], a[7], a[8]);
  const float b3 = fmaxf(a[9], a[10]);
  const uint32_t m = __float_as_uint(fmaxf(max3(b0, b1, b2), b3));
  if (__uint_as_float(m | 31u) > __uint_as_float(mk | 31u)) {
    mk = m;
    base = col0;
  }
}
}  // namespace fl

__global__ void __launch_bounds__(fl::kThreads, 1)
tc_filip_max_kernel(const __grid_constant__ CUtensorMap tx, const __grid_constant__ CUtensorMap ty,
                    const FilipFwdArgs args) {
  using namespace fl;
  const int bx = blockIdx.y;
  const int row0 = blockIdx.x * kRows;
  const int y0 = blockIdx.z * args.y_per_split, y1 = min(y0 + args.y_per_split, args.By);
  const int lx = min(args.lenX[bx], args.Tx);
  const int nh = row0 >= lx ? 0 : (row0 + kTile < lx ? 2 : 1);  // row tiles that hold at least one valid token

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* a_tiles = smem;
  uint8_t* b_tiles = a_tiles + kH * kTileBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(b_tiles + kFwdStages * kTileBytes);
  uint64_t* a_full = bars;
  uint64_t* b_full = a_full + 1;
  uint64_t* b_empty = b_full + kFwdStages;
  uint64_t* t_full = b_empty + kFwdStages;   // [kH][2]
  uint64_t* t_empty = t_full + 2 * kH;       // [kH][2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(t_empty + 2 * kH);

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;   // provably warp-uniform
  if (threadIdx.x == 0) {
    mbar_init(a_full, 1);
    for (int i = 0; i < kFwdStages; ++i) {
      mbar_init(&b_full[i], 1);
      mbar_init(&b_empty[i], 1);
    }
    for (int i = 0; i < 2 * kH; ++i) {
      mbar_init(&t_full[i], 1);
      mbar_init(&t_empty[i], 128);
    }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc<512>(tmem_slot);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0 && nh > 0) {
      tma_prefetch_desc(&tx);
      tma_prefetch_desc(&ty);
      mbar_arrive_expect_tx(a_full, (uint32_t)nh * kTileBytes);
      for (int h = 0; h < nh; ++h) {
        tma_load_3d(a_tiles + h * kTileBytes, &tx, a_full, 0, row0 + h * kTile, bx);
        tma_load_3d(a_tiles + h * kTileBytes + kPanelBytes, &tx, a_full, 64, row0 + h * kTile, bx);
      }
      int it = 0;
      for (int by = y0; by < y1; ++by) {
        const int ly = min(args.lenY[by], args.Ty);
        for (int s0 = 0; s0 < ly; s0 += kTile, ++it) {
          const int stage = it % kFwdStages;
          const uint32_t phase = (uint32_t)(it / kFwdStages) & 1u;
          mbar_wait(&b_empty[stage], phase ^ 1u);
          mbar_arrive_expect_tx(&b_full[stage], kTileBytes);
          uint8_t* dst = b_tiles + (size_t)stage * kTileBytes;
          tma_load_3d(dst, &ty, &b_full[stage], 0, s0, by);
          tma_load_3d(dst + kPanelBytes, &ty, &b_full[stage], 64, s0, by);
        }
      }
    }
  } else if (warp == 1) {
    if (nh > 0) {   // whole warp; tcgen05 instructions go out from one elected lane
      constexpr uint32_t idesc = make_idesc_bf16(kTile, kTile, 0, 0);
      mbar_wait(a_full, 0);
      int it = 0;
      for (int by = y0; by < y1; ++by) {
        const int ly = min(args.lenY[by], args.Ty);
        for (int s0 = 0; s0 < ly; s0 += kTile, ++it) {
          const int stage = it % kFwdStages;
          const uint32_t phase = (uint32_t)(it / kFwdStages) & 1u;
          const int buf = it & 1;
          const uint32_t bphase = (uint32_t)(it >> 1) & 1u;
          mbar_wait(&b_full[stage], phase);
          const uint32_t b_addr = smem_u32(b_tiles + (size_t)stage * kTileBytes);
          for (int h = 0; h < nh; ++h) {
            mbar_wait(&t_empty[h * 2 + buf], bphase ^ 1u);
            tc_fence_after_sync();
            issue_tile(tmem_base + (uint32_t)(h * 2 + buf) * kTile, smem_u32(a_tiles + h * kTileBytes), b_addr, idesc,
                       false);
            if (elect_one()) {
              umma_commit(&t_full[h * 2 + buf]);
              if (h == nh - 1) umma_commit(&b_empty[stage]);
            }
            __syncwarp();
          }
        }
      }
    }
  } else {
    const int h = (warp - 2) >> 2, q = warp & 3;
    const int t = row0 + h * kTile + q * 32 + lane;   // token row of this thread
    const bool live = h < nh;
    const int slot = blockIdx.x * (kH * 4) + h * 4 + q;
    int it = 0;
    for (int by = y0; by < y1; ++by) {
      const int ly = min(args.lenY[by], args.Ty);
      uint32_t mk = kNegKey;   // running maximum as a key: score with its 5 low mantissa bits replaced by 31 - column
      int base = 0;            // first column of the 32-column chunk the running maximum came from
      if (live) {
        for (int s0 = 0; s0 < ly; s0 += kTile, ++it) {
          const int buf = it & 1;
          const uint32_t bphase = (uint32_t)(it >> 1) & 1u;
          mbar_wait(&t_full[h * 2 + buf], bphase);
          tc_fence_after_sync();
          const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(h * 2 + buf) * kTile;
          const int nvalid = ly - s0;  // >= 1; columns >= nvalid of this tile are pads
          uint32_t ra[32], rb[32];
          if (nvalid >= kTile) {
            // full tile: the next chunk's TMEM load is in flight while the current one is reduced, and the accumulator
            // goes back to the MMA warp as soon as the last load has landed
            tmem_ld_32x32b_x32(taddr, ra);
            tmem_ld_wait();
            tmem_ld_32x32b_x32(taddr + 32, rb);
            chunk_max(ra, s0, mk, base);
            tmem_ld_wait();
            tmem_ld_32x32b_x32(taddr + 64, ra);
            chunk_max(rb, s0 + 32, mk, base);
            tmem_ld_wait();
            tmem_ld_32x32b_x32(taddr + 96, rb);
            chunk_max(ra, s0 + 64, mk, base);
            tmem_ld_wait();
            tc_fence_before_sync();
            mbar_arrive(&t_empty[h * 2 + buf]);
            chunk_max(rb, s0 + 96, mk, base);
          } else {
#pragma unroll 1
            for (int c = 0; c * 32 < nvalid; ++c) {   // warp-uniform
              tmem_ld_32x32b_x32(taddr + c * 32, ra);
              tmem_ld_wait();
              if (nvalid - c * 32 < 32) {
#pragma unroll
                for (int e = 0; e < 32; ++e)
                  if (c * 32 + e >= nvalid) ra[e] = __float_as_uint(-FLT_MAX);
              }
              chunk_max(ra, s0 + c * 32, mk, base);
            }
            tc_fence_before_sync();
            mbar_arrive(&t_empty[h * 2 + buf]);
          }
        }
      }
      const int am = base + 31 - (int)(mk & 31u);
      // the 5 bits the column index displaced are set to the middle of their range: |error| <= 2^-19 relative, unbiased
      const float m = ly > 0 ? __uint_as_float((mk & ~31u) | 16u) : -FLT_MAX;
      const int64_t pair = (int64_t)bx * args.By + by;
      if (live && t < args.Tx) args.arg[pair * args.Tx + t] = (int16_t)am;
      float v = (live && t < lx) ? m : 0.f;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane == 0) args.part[pair * args.np + slot] = v;
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
// host side
// ---------------------------------------------------------------------------------------------------------
static int np_of(int64_t T) { return (int)ceil_div(T, fl::kRows) * fl::kH * 4; }

// filip_bwd.cu
size_t filip_gather_inv_elems(int64_t Ba, int64_t Bb, int64_t TA, int64_t TB);
size_t filip_gather_grp_elems(int64_t Ba, int64_t Bb, int64_t TA, int64_t TB);
int filip_gather_side(const __nv_bfloat16* Yr, const int32_t* lenX, const int32_t* lenY, int64_t Bx, int64_t By,
                      int64_t Tx, int64_t Ty, const float* tau, const float* gX, int64_t gx0, int64_t gx1,
                      const float* gY, int64_t gy0, int64_t gy1, const int16_t* argX, const int16_t* argY, float2* w2,
                      uint32_t* inv, uint32_t* grp2, float* dX, cudaStream_t st);

// filip_tc1.cu
int tc_filip_both_fwd(const __nv_bfloat16* X, const __nv_bfloat16* Y, const int32_t* lenX, const int32_t* lenY, int64_t Bx,
                      int64_t By, int64_t Tx, int64_t Ty, const float* tau, float* partX, int npX, int16_t* argX,
                      float* partY, int16_t* argY, float* outX, int64_t sx0, int64_t sx1, float* outY, int64_t sy0,
                      int64_t sy1, cudaStream_t st);

struct FilipWs {
  __nv_bfloat16 *Ar, *Br;
  float *pa, *pb;        // forward: per-warp partial sums of the masked means
  float2* w2;            // backward (one side at a time): pair weights,
  uint32_t* inv;         //   scatter entries sorted by target row,
  uint32_t* grp2;        //   (first entry | count << 16) per group of target rows
  size_t total;
  bool ok;
};

static FilipWs filip_carve(void* ws, size_t ws_bytes, int64_t Ba, int64_t Bb, int64_t TA, int64_t TB) {
  Workspace w(ws, ws_bytes);
  FilipWs f;
  f.Ar = w.take<__nv_bfloat16>((size_t)Ba * TA * fl::kDim);
  f.Br = w.take<__nv_bfloat16>((size_t)Bb * TB * fl::kDim);
  f.pa = w.take<float>((size_t)Ba * Bb * np_of(TA));
  f.pb = w.take<float>((size_t)Ba * Bb * np_of(TB));
  f.inv = w.take<uint32_t>(filip_gather_inv_elems(Ba, Bb, TA, TB));
  f.grp2 = w.take<uint32_t>(filip_gather_grp_elems(Ba, Bb, TA, TB));
  f.w2 = w.take<float2>((size_t)Ba * Bb);
  f.total = align_up(w.off, 1024);
  f.ok = ws != nullptr && w.ok();
  return f;
}

size_t tc_filip_workspace(int64_t Ba, int64_t Bb, int64_t TA, int64_t TB) {
  return filip_carve(nullptr, 0, Ba, Bb, TA, TB).total + 1024;
}

static int filip_prep(const float* h, const int32_t* len, int64_t B, int64_t T, __nv_bfloat16* rows, cudaStream_t st) {
  const int64_t blocks = std::min<int64_t>(ceil_div(B * T, 8), (int64_t)device_sm_count() * 16);
  filip_prep_kernel<<<(unsigned)blocks, 256, 0, st>>>(h, len, B, (int)T, rows);
  PCLIP_CHECK_LAUNCH();
  return PCLIP_OK;
}

static int tc_side_fwd(const __nv_bfloat16* X, const __nv_bfloat16* Y, const int32_t* lenX, const int32_t* lenY,
                       int64_t Bx, int64_t By, int64_t Tx, int64_t Ty, const float* tau, float* part, int16_t* arg,
                       float* out, int64_t so0, int64_t so1, cudaStream_t st) {
  CUtensorMap mx, my;
  PCLIP_TRY(make_tmap_bf16_3d(&mx, X, (uint64_t)Bx, (uint64_t)Tx, fl::kDim, fl::kTile, 64));
  PCLIP_TRY(make_tmap_bf16_3d(&my, Y, (uint64_t)By, (uint64_t)Ty, fl::kDim, fl::kTile, 64));
  const int np = np_of(Tx);
  const int64_t ctas = ceil_div(Tx, fl::kRows) * Bx;
  // Split the y samples over blockIdx.z so that the grid fills whole rounds of the SMs (one CTA per SM).  The pairs of a
  // CTA are independent, so a split costs only the CTA's fixed part (TMEM allocation, barriers, its X tiles: about half a
  // sample pair).  cfg 3: 512 (row block, sample) CTAs on 148 SMs are 3.46 waves = 4 rounds of 256 pairs; two splits are
  // 6.92 waves = 7 rounds of 128 pairs, 12 % less.  Model: rounds(ys) * (pairs per CTA + 0.5); among the splits within 1.5 %
  // of the best the finest wins (measured: 8 splits cost a full-length batch 0.4 % and gain a ragged one 2 % over 2).  PCLIP_FILIP_YSPLIT overrides (A/B runs).
  const int sms = std::max(device_sm_count(), 1);
  int ysplit = 1;
  {
    const int64_t cap = std::min<int64_t>(By, std::max<int64_t>(16, ceil_div(sms, ctas)));
    auto cost = [&](int64_t ys) { return (double)ceil_div(ctas * ys, (int64_t)sms) * ((double)ceil_div(By, ys) + 0.5); };
    double best = 1e300;
    for (int64_t ys = 1; ys <= cap; ++ys) best = std::min(best, cost(ys));
    for (int64_t ys = 1; ys <= cap; ++ys)
      if (cost(ys) <= best * 1.015) ysplit = (int)ys;      // the finest split within 1.5 % of the best
    static const int forced = [] {
      const char* e = getenv("PCLIP_FILIP_YSPLIT");
      return e ? atoi(e) : 0;
    }();
    if (forced >= 1) ysplit = (int)std::min<int64_t>(forced, By);
  }
  const int yps = (int)ceil_div(By, ysplit);
  ysplit = (int)ceil_div(By, yps);
  FilipFwdArgs a{(int)Bx, (int)By, (int)Tx, (int)Ty, lenX, lenY, yps, part, np, arg};
  PCLIP_CHECK_CUDA(
      cudaFuncSetAttribute(tc_filip_max_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fl::kFwdSmem));
  dim3 grid((unsigned)ceil_div(Tx, fl::kRows), (unsigned)Bx, (unsigned)ysplit);
  tc_filip_max_kernel<<<grid, fl::kThreads, fl::kFwdSmem, st>>>(mx, my, a);
  PCLIP_CHECK_LAUNCH();
  return filip_finalize(part, np, lenX, (int)Bx, (int)By, (int)Tx, tau, out, so0, so1, st);
}

int tc_filip_fwd(const float* hA, const float* hB, const int32_t* lenA, const int32_t* lenB, int64_t Ba, int64_t Bb,
                 int64_t TA, int64_t TB, int D, const float* tau, float* simA, float* simB, int16_t* argA,
                 int16_t* argB, void* ws, size_t ws_bytes, cudaStream_t st) {
  PCLIP_REQUIRE(D == fl::kDim, PCLIP_ERR_UNSUPPORTED, "filip: bf16 tensor-core path needs D == %d (got %d)", fl::kDim, D);
  PCLIP_REQUIRE((reinterpret_cast<uintptr_t>(hA) & 15) == 0 && (reinterpret_cast<uintptr_t>(hB) & 15) == 0,
                PCLIP_ERR_INVALID, "filip: hA / hB must be 16-byte aligned");
  FilipWs f = filip_carve(ws, ws_bytes, Ba, Bb, TA, TB);
  PCLIP_REQUIRE(f.ok, PCLIP_ERR_WORKSPACE, "filip_fwd: workspace too small");
  PCLIP_TRY(filip_prep(hA, lenA, Ba, TA, f.Ar, st));
  PCLIP_TRY(filip_prep(hB, lenB, Bb, TB, f.Br, st));
  // One sweep for both sides (filip_tc1.cu) when one modality's samples fit a CTA's 4 stationary row tiles; otherwise
  // (or with PCLIP_FILIP_FWD=2, for A/B runs) one launch per side.
  static const int mode = [] {
    const char* e = getenv("PCLIP_FILIP_FWD");
    return e ? atoi(e) : 0;
  }();
  if (mode != 2 && TA <= 512)
    return tc_filip_both_fwd(f.Ar, f.Br, lenA, lenB, Ba, Bb, TA, TB, tau, f.pa, np_of(TA), argA, f.pb, argB, simA, Bb, 1,
                             simB, 1, Bb, st);
  if (mode != 2 && TB <= 512)
    return tc_filip_both_fwd(f.Br, f.Ar, lenB, lenA, Bb, Ba, TB, TA, tau, f.pb, np_of(TB), argB, f.pa, argA, simB, 1, Bb,
                             simA, Bb, 1, st);
  PCLIP_TRY(tc_side_fwd(f.Ar, f.Br, lenA, lenB, Ba, Bb, TA, TB, tau, f.pa, argA, simA, Bb, 1, st));
  PCLIP_TRY(tc_side_fwd(f.Br, f.Ar, lenB, lenA, Bb, Ba, TB, TA, tau, f.pb, argB, simB, 1, Bb, st));
  return PCLIP_OK;
}

int tc_filip_bwd(const float* hA, const float* hB, const int32_t* lenA, const int32_t* lenB, int64_t Ba, int64_t Bb,
                 int64_t TA, int64_t TB, int D, const float* tau, const int16_t* argA, const int16_t* argB,
                 const float* gA, const float* gB, float* dhA, float* dhB, void* ws, size_t ws_bytes, cudaStream_t st) {
  PCLIP_REQUIRE(D == fl::kDim, PCLIP_ERR_UNSUPPORTED, "filip: bf16 tensor-core path needs D == %d (got %d)", fl::kDim, D);
  PCLIP_REQUIRE((reinterpret_cast<uintptr_t>(hA) & 15) == 0 && (reinterpret_cast<uintptr_t>(hB) & 15) == 0,
                PCLIP_ERR_INVALID, "filip: hA / hB must be 16-byte aligned");
  FilipWs f = filip_carve(ws, ws_bytes, Ba, Bb, TA, TB);
  PCLIP_REQUIRE(f.ok, PCLIP_ERR_WORKSPACE, "filip_bwd: workspace too small");
  // the bf16 token rows are rebuilt (HBM-bound, ~1% of the step) so that fwd and bwd may use different workspaces
  PCLIP_TRY(filip_prep(hA, lenA, Ba, TA, f.Ar, st));
  PCLIP_TRY(filip_prep(hB, lenB, Bb, TB, f.Br, st));
  // dhA: gather through argA, scatter through the inverse of argB; then the mirror image for dhB
  PCLIP_TRY(filip_gather_side(f.Br, lenA, lenB, Ba, Bb, TA, TB, tau, gA, Bb, 1, gB, 1, Bb, argA, argB, f.w2, f.inv, f.grp2,
                              dhA, st));
  PCLIP_TRY(filip_gather_side(f.Ar, lenB, lenA, Bb, Ba, TB, TA, tau, gB, 1, Bb, gA, Bb, 1, argB, argA, f.w2, f.inv, f.grp2,
                              dhB, st));
  return PCLIP_OK;
}

}  // namespace pclip
