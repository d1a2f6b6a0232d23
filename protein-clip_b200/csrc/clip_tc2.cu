// BF16 tensor-core kernels of the CLIP InfoNCE contractions at configuration size (PCLIP_COMPUTE_BF16, D == 128): the
// single-pass forward (tc2_sumexp_kernel) and the transposed-form gradient (tc3_grad_kernel).
//   ref: modules/training_utils.py:85-89 (S = x @ y.T; exp; row / column sums; loss) and its autograd backward.
//
// Differences from the two-sided forward in clip_tc.cu (the path below 2048 columns):
//   * 256-row CTAs: two 128-row UMMA tiles share every streamed column tile (half the L2 -> SMEM operand traffic).
//   * the column tile is loaded ONCE: the second MMA of the gradient (P . Y) reads the same shared-memory tile as an
//     MN-major B operand, so no transposed copy of the embeddings exists in HBM any more.
//   * 8 epilogue warps (two per SM sub-partition), tcgen05.ld software-pipelined one 32-column chunk ahead.
//   * exp2 is split between the MUFU (ex2.approx) and a Cody-Waite + minimax polynomial on the FMA pipes (packed
//     f32x2 FFMA2), because at D = 128 the MUFU rate (16 / clk / SM) -- not the tensor core -- bounds a score tile.
//   * tc2_sumexp_kernel is single-pass: row sums in-thread, column sums by a third, M = 64 "ones" MMA over the bf16
//     exp tile (the tile is the MN-major B operand), instead of recomputing the transposed score matrix.
#include <cuda_bf16.h>
#include <stdio.h>
#include <stdlib.h>

#include "common.cuh"
#include "ptx.cuh"

#ifndef PCLIP_FWD_POLY
#define PCLIP_FWD_POLY 10
#endif
#ifndef PCLIP_BWD_POLY
#define PCLIP_BWD_POLY 8
#endif

namespace pclip {
namespace c2 {

constexpr int kTile = 128;
constexpr int kDim = 128;
constexpr uint32_t kPanelBytes = 128 * 128;
constexpr uint32_t kTileBytes = 2 * kPanelBytes;
constexpr int kH = 2;
constexpr int kRows = kH * kTile;
constexpr int kThreads = 64 + 32 * 4 * kH;
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kMagic = 12582912.0f;  // 1.5 * 2^23: (x + kMagic) holds round(x) in its low mantissa bits

// Operand descriptors are built once per tile (`*_base`) and advanced per k-step by a constant (`*_step`).
// K-major operand ([rows][64 bf16] panels): k-step k covers 16 bf16 = 32 B inside a 128 B swizzled row.
__device__ __forceinline__ uint64_t kmajor_base(uint32_t tile_addr) { return make_smem_desc_sw128(tile_addr, 16, 1024); }
__device__ __forceinline__ uint64_t kmajor_step(uint64_t base, int k) {
  return desc_advance(base, (uint32_t)(k >> 2) * kPanelBytes + (uint32_t)(k & 3) * 32u);
}
// The same tile read as an MN-major operand (MN runs along the 128 B rows, K along the tile's rows): k-step k covers
// tile rows [16k, 16k+16) = two 8-row swizzle atoms (SBO = 1024 B); the two 64-element MN halves are the two panels
// (LBO = kPanelBytes).
__device__ __forceinline__ uint64_t mnmajor_base(uint32_t tile_addr) {
  return make_smem_desc_sw128(tile_addr, kPanelBytes, 1024);
}
__device__ __forceinline__ uint64_t mnmajor_step(uint64_t base, int k) { return desc_advance(base, (uint32_t)k * 2048u); }

__device__ __forceinline__ uint64_t pack2(float lo, float hi) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void unpack2(uint64_t v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ uint64_t fma2(uint64_t a, uint64_t b, uint64_t c) {
  uint64_t d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ uint64_t add2(uint64_t a, uint64_t b) {
  uint64_t d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ uint64_t sub2(uint64_t a, uint64_t b) {
  uint64_t d;
  asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ uint64_t mul2(uint64_t a, uint64_t b) {
  uint64_t d;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}

// exp(x) for a pair on the FMA pipes: y = x log2(e) = n + f, |f| <= 1/2; 2^f by a minimax polynomial (relative error
// 7.5e-5 at degree 3, 2.6e-6 at degree 4), 2^n by adding n to the exponent field.  |y| must stay below 126.
template <int kDegree>
__device__ __forceinline__ void exp_poly2(float x0, float x1, float& e0, float& e1) {
  const uint64_t x = pack2(x0, x1);
  const uint64_t l2 = pack2(kLog2e, kLog2e);
  const uint64_t t = fma2(x, l2, pack2(kMagic, kMagic));
  const uint64_t nn = sub2(pack2(kMagic, kMagic), t);   // -n
  const uint64_t f = fma2(x, l2, nn);
  uint64_t p;
  if (kDegree == 3) {
    p = fma2(f, pack2(0.055171459913253784f, 0.055171459913253784f), pack2(0.2426108568906784f, 0.2426108568906784f));
    p = fma2(p, f, pack2(0.6932609677314758f, 0.6932609677314758f));
    p = fma2(p, f, pack2(0.9999281167984009f, 0.9999281167984009f));
  } else {
    p = fma2(f, pack2(0.009570082649588585f, 0.009570082649588585f), pack2(0.05591784417629242f, 0.05591784417629242f));
    p = fma2(p, f, pack2(0.240247443318367f, 0.240247443318367f));
    p = fma2(p, f, pack2(0.6931217908859253f, 0.6931217908859253f));
    p = fma2(p, f, pack2(0.9999992847442627f, 0.9999992847442627f));
  }
  float p0, p1, t0, t1;
  unpack2(p, p0, p1);
  unpack2(t, t0, t1);
  e0 = __int_as_float(__float_as_int(p0) + (__float_as_int(t0) << 23));
  e1 = __int_as_float(__float_as_int(p1) + (__float_as_int(t1) << 23));
}

__device__ __forceinline__ uint32_t bf16x2_bits(float lo, float hi) {
  __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);  // .x (low half) = lo = the lower column
  return *reinterpret_cast<uint32_t*>(&h);
}

}  // namespace c2

// =================================================================================================================
// gradient: da_i = scale * ( sum_j exp(a_i . b_j) (rinv_i + cinv_j) b_j  -  2 b_(off + i) )
// =================================================================================================================
// debug timeline (PCLIP_FWD_TRACE / PCLIP_BWD_TRACE = file): SM clocks of pipeline events of CTA (0,0,0)
namespace c2 {
constexpr int kTraceItems = 96;
}
#define PCLIP_TRACE(ev, u)                                                                       \
  do {                                                                                          \
    if (tr && (u) < kTraceItems) tr[(ev) * kTraceItems + (u)] = clock64();                      \
  } while (0)

struct Tc2BwdSide {
  int64_t Ma, Nb, diag_offset;
  const float* rinv;
  const float* cinv;
  const __nv_bfloat16* b;
  float* out;  // direct: da [Ma,128]; split: partial [nsplit][Ma][128]
};
struct Tc2BwdArgs {
  Tc2BwdSide s[2];
  int tiles_per_split;
  int direct;
  float scale;
  const float* gscale;
  long long* trace;
};

namespace c2 {
constexpr int kPolyPerChunk = PCLIP_BWD_POLY;  // of every 32 exps, this many go to the FMA pipes (multiple of 2)
}  // namespace c2

// =================================================================================================================
// gradient, transposed formulation (third generation): both contractions are issued with N = 256.
//   S^T [128 cols j x 256 rows i] = Y_t . A^T          (A = column tile, K-major;  B = the CTA's 256 rows, K-major)
//   dA^T[128 d      x 256 rows i] += Y_t^T . P^T        (A = column tile read MN-major; B = P^T [j][i], MN-major)
// Why: a tcgen05.mma with both operands in shared memory is paced by the operand read (~96 B/clk): M128 N128 K16 takes
// 86 clk, N256 takes 128 clk (tools/probe_mma.cu), i.e. N = 256 is the only shape that runs at the tensor-pipe rate,
// and it halves the SMEM operand traffic per score.  The epilogue thread owns one COLUMN j (TMEM lane): 1/C_j lives in
// a register, 1/R_i of the 256 rows is staged in SMEM once per CTA.  S^T has one accumulator (TMEM is full with dA^T):
// it is released as soon as the epilogue holds the scores in registers, so S^T(t+1) runs under the epilogue of tile t.
// =================================================================================================================
#ifndef PCLIP_RINV_LD
#define PCLIP_RINV_LD 128
#endif
namespace c3 {
using namespace c2;
constexpr int kStages = 3;
constexpr int kEpiWarps = 16;
constexpr int kThreads3 = 64 + 32 * kEpiWarps;
constexpr uint32_t kABytes = kRows * kDim * 2;       // 64 KB, panel-major: [2 panels][256 rows][64 d]
constexpr uint32_t kPBytes = kTile * kRows * 2;      // 64 KB: [4 panels of 64 rows i][128 j][64 i]
// A | Y ring | P^T | rinv staging [256] | barriers
constexpr uint32_t kSmem3 = kABytes + kStages * kTileBytes + kPBytes + kRows * 4 + 256 + 768;
static_assert(kSmem3 <= 232448, "shared memory budget");
}  // namespace c3

__global__ void __launch_bounds__(c3::kThreads3, 1)
tc3_grad_kernel(const __grid_constant__ CUtensorMap ta0, const __grid_constant__ CUtensorMap tb0,
                const __grid_constant__ CUtensorMap ta1, const __grid_constant__ CUtensorMap tb1,
                const Tc2BwdArgs args) {
  using namespace c3;
  const Tc2BwdSide& sd = args.s[blockIdx.z];
  const CUtensorMap* ta = blockIdx.z == 0 ? &ta0 : &ta1;
  const CUtensorMap* tb = blockIdx.z == 0 ? &tb0 : &tb1;
  const int64_t row0 = (int64_t)blockIdx.x * kRows;
  if (row0 >= sd.Ma) return;
  const int ntiles_all = (int)((sd.Nb + kTile - 1) / kTile);
  const int tile_begin = blockIdx.y * args.tiles_per_split;
  const int ntiles = max(min(tile_begin + args.tiles_per_split, ntiles_all) - tile_begin, 0);

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* a_rows = smem;
  uint8_t* y_tiles = a_rows + kABytes;
  uint8_t* p_t = y_tiles + kStages * kTileBytes;
  float* rinv_s = reinterpret_cast<float*>(p_t + kPBytes);
  uint64_t* bars = reinterpret_cast<uint64_t*>(rinv_s + kRows);
  uint64_t* a_full = bars;                 // 1
  uint64_t* y_full = a_full + 1;           // kStages
  uint64_t* y_empty = y_full + kStages;    // kStages
  uint64_t* s_full = y_empty + kStages;    // 1: S^T(t) is in TMEM
  uint64_t* s_free = s_full + 1;           // 1: every epilogue thread holds its scores of tile t in registers
  uint64_t* p_full = s_free + 1;           // 1: P^T(t) is in SMEM
  uint64_t* pv_done = p_full + 1;          // 1: the second MMA has consumed P^T(t)
  uint64_t* d_full = pv_done + 1;          // 1
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(d_full + 1);
  if (threadIdx.x == 0 && reinterpret_cast<uint8_t*>(tmem_slot + 1) > smem_raw + kSmem3) __trap();  // alignment slack

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;   // provably warp-uniform
  long long* tr = (args.trace && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0) ? args.trace : nullptr;
  if (threadIdx.x == 0) {
    mbar_init(a_full, 1);
    for (int i = 0; i < kStages; ++i) {
      mbar_init(&y_full[i], 1);
      mbar_init(&y_empty[i], 1);
    }
    mbar_init(s_full, 1);
    mbar_init(s_free, 32 * kEpiWarps);
    mbar_init(p_full, 32 * kEpiWarps);
    mbar_init(pv_done, 1);
    mbar_init(d_full, 1);
    fence_mbar_init();
  }
  if (threadIdx.x < kRows) {
    const int64_t r = row0 + threadIdx.x;
    rinv_s[threadIdx.x] = r < sd.Ma ? sd.rinv[r] : 0.f;
  }
  if (warp == 1) tmem_alloc<512>(tmem_slot);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_s = *tmem_slot;            // columns [0, 256): S^T; [256, 512): dA^T
  const uint32_t tmem_d = tmem_s + kRows;

  if (warp == 0) {
    if (lane == 0) {
      tma_prefetch_desc(ta);
      tma_prefetch_desc(tb);
      mbar_arrive_expect_tx(a_full, kABytes);    // rows beyond Ma are zero-filled by TMA (the bytes still count)
      for (int p = 0; p < 2; ++p)
        for (int h = 0; h < kH; ++h)
          tma_load_2d(a_rows + p * (kABytes / 2) + h * kPanelBytes, ta, a_full, p * 64, (int32_t)(row0 + h * kTile));
      for (int t = 0; t < ntiles; ++t) {
        const int stage = t % kStages;
        mbar_wait(&y_empty[stage], ((uint32_t)(t / kStages) & 1u) ^ 1u);
        mbar_arrive_expect_tx(&y_full[stage], kTileBytes);
        const int32_t col0 = (tile_begin + t) * kTile;
        uint8_t* dst = y_tiles + (size_t)stage * kTileBytes;
        tma_load_2d(dst, tb, &y_full[stage], 0, col0);
        tma_load_2d(dst + kPanelBytes, tb, &y_full[stage], 64, col0);
      }
    }
  } else if (warp == 1) {
    if (ntiles > 0) {   // the whole warp runs the issue loop; one elected lane issues each tcgen05 instruction
      constexpr uint32_t idesc_s = make_idesc_bf16(kTile, kRows, 0, 0);   // S^T = Y . A^T      (both K-major)
      constexpr uint32_t idesc_g = make_idesc_bf16(kTile, kRows, 1, 1);   // dA^T += Y^T . P^T  (both MN-major)
      const uint64_t ad = kmajor_base(smem_u32(a_rows)), pd = mnmajor_base(smem_u32(p_t));
      mbar_wait(a_full, 0);
      auto issue_s = [&](int t) {
        const int stage = t % kStages;
        mbar_wait(&y_full[stage], (uint32_t)(t / kStages) & 1u);
        if (t > 0) mbar_wait(s_free, (uint32_t)(t - 1) & 1u);
        tc_fence_after_sync();
        if (lane == 0) PCLIP_TRACE(0, t);
        const uint64_t yd = kmajor_base(smem_u32(y_tiles + (size_t)stage * kTileBytes));
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < 8; ++k)
            umma_bf16_ss(tmem_s, kmajor_step(yd, k), desc_advance(ad, (uint32_t)(k >> 2) * (kABytes / 2) + (uint32_t)(k & 3) * 32u),
                         idesc_s, k > 0 ? 1u : 0u);
          umma_commit(s_full);
        }
        __syncwarp();
      };
      issue_s(0);
      for (int t = 0; t < ntiles; ++t) {
        if (t + 1 < ntiles) issue_s(t + 1);
        mbar_wait(p_full, (uint32_t)t & 1u);
        tc_fence_after_sync();
        if (lane == 0) PCLIP_TRACE(1, t);
        const uint64_t yd = mnmajor_base(smem_u32(y_tiles + (size_t)(t % kStages) * kTileBytes));
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < 8; ++k)
            umma_bf16_ss(tmem_d, mnmajor_step(yd, k), mnmajor_step(pd, k), idesc_g, (t > 0 || k > 0) ? 1u : 0u);
          umma_commit(pv_done);
          umma_commit(&y_empty[t % kStages]);
          if (t == ntiles - 1) umma_commit(d_full);
        }
        __syncwarp();
      }
    }
  } else {
    // 16 epilogue warps: lane quadrant q (columns j of the tile) x row group g (64 rows i each)
    const int g = (warp - 2) >> 2, q = warp & 3;
    const int j = q * 32 + lane;                         // my column inside the tile == my TMEM lane
    const uint32_t lane_base = (uint32_t)(q * 32) << 16;
    const uint32_t p_row = smem_u32(p_t) + (uint32_t)g * kPanelBytes;   // panel g holds rows i in [64 g, 64 g + 64)
    const bool tracer = threadIdx.x == 64;
    float cinv = (int64_t)tile_begin * kTile + j < sd.Nb && ntiles > 0 ? sd.cinv[(int64_t)tile_begin * kTile + j] : 0.f;
    for (int t = 0; t < ntiles; ++t) {
      const uint64_t cinv2 = pack2(cinv, cinv);
      const int64_t jn = (int64_t)(tile_begin + t + 1) * kTile + j;       // prefetch the next tile's 1/C_j
      const float cinv_next = (t + 1 < ntiles && jn < sd.Nb) ? sd.cinv[jn] : 0.f;
      mbar_wait(s_full, (uint32_t)t & 1u);
      tc_fence_after_sync();
      if (tracer) PCLIP_TRACE(2, t);
      uint32_t packed[2][16];
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        uint32_t v[32];
        tmem_ld_32x32b_x32(tmem_s + lane_base + (uint32_t)(g * 64 + c * 32), v);
        tmem_ld_wait();
        if (c == 1) {                          // all my scores are in registers: S^T(t+1) may overwrite the accumulator
          tc_fence_before_sync();
          mbar_arrive(s_free);
          if (tracer) PCLIP_TRACE(3, t);
        }
        // explicit shared-space loads: through a generic pointer these were LD.E.128 at 6.3 wavefronts each (r01c)
        const uint32_t rs_addr = smem_u32(rinv_s + g * 64 + c * 32);
        uint64_t rs2[16];
#pragma unroll
        for (int e = 0; e < 32; e += PCLIP_RINV_LD / 32) {
          if (PCLIP_RINV_LD == 128)
            asm volatile("ld.shared.v2.b64 {%0, %1}, [%2];" : "=l"(rs2[e >> 1]), "=l"(rs2[(e >> 1) + 1]) : "r"(rs_addr + 4u * e) : "memory");
          else
            asm volatile("ld.shared.b64 %0, [%1];" : "=l"(rs2[e >> 1]) : "r"(rs_addr + 4u * e) : "memory");
        }
#pragma unroll
        for (int e = 0; e < 32; e += 2) {
          float e0, e1;
          if (e < kPolyPerChunk) {
            exp_poly2<3>(__uint_as_float(v[e]), __uint_as_float(v[e + 1]), e0, e1);
          } else {
            float y0, y1;
            unpack2(mul2(pack2(__uint_as_float(v[e]), __uint_as_float(v[e + 1])), pack2(kLog2e, kLog2e)), y0, y1);
            e0 = ex2_approx(y0);
            e1 = ex2_approx(y1);
          }
          const uint64_t w = add2(cinv2, rs2[e >> 1]);
          float p0, p1;
          unpack2(mul2(pack2(e0, e1), w), p0, p1);
          packed[c][e >> 1] = bf16x2_bits(p0, p1);
        }
      }
      if (tracer) PCLIP_TRACE(4, t);
      if (t > 0) mbar_wait(pv_done, (uint32_t)(t - 1) & 1u);   // P^T(t-1) has been consumed by the second MMA
      if (tracer) PCLIP_TRACE(5, t);
#pragma unroll
      for (int c = 0; c < 2; ++c)
#pragma unroll
        for (int gg = 0; gg < 4; ++gg)
          st_shared_v4(p_row + sw128_offset((uint32_t)j, (uint32_t)(c * 4 + gg)), packed[c][4 * gg], packed[c][4 * gg + 1],
                       packed[c][4 * gg + 2], packed[c][4 * gg + 3]);
      fence_proxy_async_smem();
      mbar_arrive(p_full);
      if (tracer) PCLIP_TRACE(6, t);
      cinv = cinv_next;
    }
    if (ntiles > 0) {
      mbar_wait(d_full, 0);
      tc_fence_after_sync();
    }
    // ---- final: dA^T accumulator (lane = d, column = row i) -> global; a warp store covers 32 consecutive d ----
    const float s = args.scale * (args.gscale ? args.gscale[0] : 1.0f);
    const int d = j;
#pragma unroll 1
    for (int c = 0; c < 2; ++c) {
      uint32_t r[32];
      if (ntiles > 0) {
        tmem_ld_32x32b_x32(tmem_d + lane_base + (uint32_t)(g * 64 + c * 32), r);
        tmem_ld_wait();
      } else {
#pragma unroll
        for (int e = 0; e < 32; ++e) r[e] = 0u;
      }
#pragma unroll
      for (int e = 0; e < 32; ++e) {
        const int64_t row = row0 + g * 64 + c * 32 + e;
        if (row < sd.Ma) {
          if (args.direct) {
            const float bpos = __bfloat162float(sd.b[(sd.diag_offset + row) * kDim + d]);
            sd.out[row * kDim + d] = s * (__uint_as_float(r[e]) - 2.0f * bpos);
          } else {
            sd.out[((int64_t)blockIdx.y * sd.Ma + row) * kDim + d] = __uint_as_float(r[e]);
          }
        }
      }
    }
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after_sync();
    __syncwarp();
    tmem_dealloc<512>(tmem_s);
  }
}

// =================================================================================================================
// single-pass sum-exp: rowsum_i = sum_j exp(a_i . b_j) (fp32 exps), colsum_j = sum_{i != pos(j)} bf16(exp(a_i . b_j))
// over THIS launch's rows (the diagonal term is added back in fp32 by col_reduce_kernel), diag_i = a_i . b_(off+i).
//
// CTA = 256 rows (two 128-row A tiles share every streamed column tile: the kernel was bound by the L2 -> SMEM
// latency of the column tiles with 128-row CTAs, profiles/r01c).  Work items u = (column tile t, row half h).
//   warp 0      TMA producer (A tiles once, column-tile ring)
//   warp 1      score-MMA issuer: S(u) = A_h Y_t^T into one of three TMEM accumulators, up to three items ahead
//   warps 2-17  epilogue, eight warps per row half (4 lane quadrants x 2 column halves): thread = one row x 64 columns
//               of its half's item, two 32-column TMEM loads; the two halves' items are processed side by side
//   warp 18     column-sum MMA issuer: C_t += E(u)^T . ones (M = 128 columns, N = 16, the bf16 exp tile is the
//               MN-major A operand) as soon as item u's exp tile is complete -- in one issue loop with the score
//               MMAs each waited on the other's barriers
// =================================================================================================================
struct Tc2FwdArgs {
  int64_t Ma, Nb, diag_offset;
  float* rowpart;   // [4 * nsplit][Ma]  (four column quarters per split)
  float* colpart;   // [256-row blocks][Nb]
  float* diag;      // [Ma]
  int tiles_per_split;
  long long* trace; // debug timeline of CTA (0,0): [8][kTraceItems] SM clocks, or NULL (PCLIP_FWD_TRACE=file)
};

#ifndef PCLIP_FWD_STAGES
#define PCLIP_FWD_STAGES 2
#define PCLIP_FWD_EBUF 3
#endif
namespace c2 {
constexpr int kFwdStages = PCLIP_FWD_STAGES;   // column-tile ring
constexpr int kEBuf = PCLIP_FWD_EBUF;          // bf16 exp tiles awaiting their column-sum MMA, indexed by item % kEBuf
constexpr int kSBuf = 3;                    // score accumulators in TMEM
constexpr uint32_t kOnesBytes = 2048;       // B operand of the column-sum MMA: [16 x 64] bf16 ones (one K panel, reused)
// A x2 | Y ring | E x2 | ones | barriers
constexpr uint32_t kFwdSmem = kTileBytes * (kH + kFwdStages + kEBuf) + kOnesBytes + 256 + 768;
static_assert(kFwdSmem <= 232448, "shared memory budget");
constexpr int kFwdPoly = PCLIP_FWD_POLY;  // of every 32 exps, this many on the FMA pipes (degree 4)
constexpr int kFwdEpiWarps = 16;          // 4 warps per SM sub-partition
constexpr int kFwdThreads = 64 + 32 * kFwdEpiWarps + 32;   // + the column-sum issuer warp
}  // namespace c2

template <int kPoly>
__global__ void __launch_bounds__(c2::kFwdThreads, 1)
tc2_sumexp_kernel(const __grid_constant__ CUtensorMap ta, const __grid_constant__ CUtensorMap tb, const Tc2FwdArgs args) {
  using namespace c2;
  const int64_t row0 = (int64_t)blockIdx.x * kRows;
  const int nh = row0 + kTile < args.Ma ? 2 : 1;
  const int ntiles_all = (int)((args.Nb + kTile - 1) / kTile);
  const int tile_begin = blockIdx.y * args.tiles_per_split;
  const int ntiles = max(min(tile_begin + args.tiles_per_split, ntiles_all) - tile_begin, 0);

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* a_tiles = smem;                                  // [kH]
  uint8_t* y_tiles = a_tiles + kH * kTileBytes;             // [kFwdStages]
  uint8_t* e_tiles = y_tiles + kFwdStages * kTileBytes;     // [kEBuf]
  uint8_t* ones = e_tiles + kEBuf * kTileBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(ones + kOnesBytes);
  uint64_t* a_full = bars;                   // 1
  uint64_t* y_full = a_full + 1;             // kFwdStages
  uint64_t* y_empty = y_full + kFwdStages;   // kFwdStages
  uint64_t* s_full = y_empty + kFwdStages;   // kSBuf
  uint64_t* s_empty = s_full + kSBuf;        // kSBuf
  uint64_t* e_full = s_empty + kSBuf;        // kEBuf
  uint64_t* e_empty = e_full + kEBuf;        // kEBuf
  uint64_t* c_full = e_empty + kEBuf;        // 2
  uint64_t* c_empty = c_full + 2;            // 2
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(c_empty + 2);
  if (threadIdx.x == 0 && reinterpret_cast<uint8_t*>(tmem_slot + 1) > smem_raw + kFwdSmem) __trap();  // alignment slack

  // the warp index through a shuffle: provably warp-uniform, so the role branches and the MMA warp's loop state are
  // uniform (uniform registers feed tcgen05.mma directly instead of an elect / R2UR loop per operand)
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;
  long long* tr = (args.trace && blockIdx.x == 0 && blockIdx.y == 0) ? args.trace : nullptr;
  if (threadIdx.x == 0) {
    mbar_init(a_full, 1);
    for (int i = 0; i < kFwdStages; ++i) {
      mbar_init(&y_full[i], 1);
      mbar_init(&y_empty[i], 1);
    }
    for (int i = 0; i < kSBuf; ++i) {
      mbar_init(&s_full[i], 1);
      mbar_init(&s_empty[i], 16 * kFwdEpiWarps);   // the eight warps of the item's row half
    }
    for (int i = 0; i < kEBuf; ++i) {
      mbar_init(&e_full[i], 16 * kFwdEpiWarps);
      mbar_init(&e_empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&c_full[i], 1);
      mbar_init(&c_empty[i], 128);
    }
    fence_mbar_init();
  }
  for (uint32_t o = threadIdx.x * 4u; o < kOnesBytes; o += kFwdThreads * 4u) *reinterpret_cast<uint32_t*>(ones + o) = 0x3F803F80u;
  fence_proxy_async_smem();
  if (warp == 1) tmem_alloc<512>(tmem_slot);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;     // [0, 384): three S accumulators; [384, 416): column sums, 16 columns x 2
  const uint32_t tmem_c = tmem_base + kSBuf * kTile;

  if (warp == 0) {
    if (lane == 0) {
      tma_prefetch_desc(&ta);
      tma_prefetch_desc(&tb);
      mbar_arrive_expect_tx(a_full, (uint32_t)nh * kTileBytes);
      for (int h = 0; h < nh; ++h) {
        tma_load_2d(a_tiles + h * kTileBytes, &ta, a_full, 0, (int32_t)(row0 + h * kTile));
        tma_load_2d(a_tiles + h * kTileBytes + kPanelBytes, &ta, a_full, 64, (int32_t)(row0 + h * kTile));
      }
      for (int t = 0; t < ntiles; ++t) {
        const int stage = t % kFwdStages;
        mbar_wait(&y_empty[stage], ((uint32_t)(t / kFwdStages) & 1u) ^ 1u);
        mbar_arrive_expect_tx(&y_full[stage], kTileBytes);
        const int32_t col0 = (tile_begin + t) * kTile;
        uint8_t* dst = y_tiles + (size_t)stage * kTileBytes;
        tma_load_2d(dst, &tb, &y_full[stage], 0, col0);
        tma_load_2d(dst + kPanelBytes, &tb, &y_full[stage], 64, col0);
      }
    }
  } else if (warp == 1) {
    if (ntiles > 0) {   // the whole warp runs the issue loop; one elected lane issues each tcgen05 instruction
      constexpr uint32_t idesc_s = make_idesc_bf16(kTile, kTile, 0, 0);  // S = A . Y^T
      const int n_items = ntiles * nh;
      mbar_wait(a_full, 0);
      // Score MMAs run up to kSBuf items ahead of the epilogue, with blocking waits (a polling loop lost ~1000 clk per
      // item: the issuing thread shares its scheduler with four epilogue warps and was starved while they were busy).
      auto issue_s = [&](int u) {
        const int t = u / nh, h = u - t * nh;
        const int sb = u % kSBuf, stage = t % kFwdStages;
        mbar_wait(&s_empty[sb], ((uint32_t)(u / kSBuf) & 1u) ^ 1u);
        if (h == 0) mbar_wait(&y_full[stage], (uint32_t)(t / kFwdStages) & 1u);
        tc_fence_after_sync();
        if (lane == 0) PCLIP_TRACE(1, u);
        const uint64_t yd = kmajor_base(smem_u32(y_tiles + (size_t)stage * kTileBytes));
        const uint64_t ad = kmajor_base(smem_u32(a_tiles + (size_t)h * kTileBytes));
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < 8; ++k)
            umma_bf16_ss(tmem_base + (uint32_t)sb * kTile, kmajor_step(ad, k), kmajor_step(yd, k), idesc_s, k > 0 ? 1u : 0u);
          if (h == nh - 1) umma_commit(&y_empty[stage]);
          umma_commit(&s_full[sb]);
        }
        __syncwarp();
        if (lane == 0) PCLIP_TRACE(7, u);
      };
      for (int u = 0; u < n_items; ++u) issue_s(u);
    }
  } else if (warp == 2 + kFwdEpiWarps) {
    if (ntiles > 0) {   // column-sum issuer: its own warp, so that a wait for an exp tile never holds back the score MMAs
      constexpr uint32_t idesc_c = make_idesc_bf16(kTile, 16, 1, 0);     // C[col, :] = sum_rows E[row, col] (E^T . ones)
      const uint64_t od = kmajor_base(smem_u32(ones));
      const int n_items = ntiles * nh;
      for (int u = 0; u < n_items; ++u) {
        const int t = u / nh, h = u - t * nh;
        const int eb = u % kEBuf;
        mbar_wait(&e_full[eb], (uint32_t)(u / kEBuf) & 1u);
        if (h == 0) mbar_wait(&c_empty[t & 1], ((uint32_t)(t >> 1) & 1u) ^ 1u);
        tc_fence_after_sync();
        if (lane == 0) PCLIP_TRACE(2, u);
        const uint64_t ed = mnmajor_base(smem_u32(e_tiles + (size_t)eb * kTileBytes));
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < 8; ++k)
            umma_bf16_ss(tmem_c + (uint32_t)(t & 1) * 16u, mnmajor_step(ed, k), desc_advance(od, (uint32_t)(k & 3) * 32u),
                         idesc_c, (h > 0 || k > 0) ? 1u : 0u);
          umma_commit(&e_empty[eb]);
          if (h == nh - 1) umma_commit(&c_full[t & 1]);
        }
        __syncwarp();
        if (lane == 0) PCLIP_TRACE(0, u);
      }
    }
  } else if (warp < 2 + kFwdEpiWarps) {
    // Eight warps per row half: warp = (row half, 64-column half, lane quadrant); thread = one row x 64 columns of its
    // half's item, read as two 32-column TMEM loads.  The two halves' items are in flight side by side, and the
    // per-item barrier / store / fence overhead is paid once per 64 columns instead of once per 32.
    const int w = warp - 2;
    const int grp = w >> 3;                  // row half this warp serves: items u = t * nh + grp
    const int colhalf = (w >> 2) & 1;        // columns [64 * colhalf, +64) of the item == K panel of the exp tile
    const int q = warp & 3;                  // TMEM lane quadrant this warp may read
    const int i = q * 32 + lane;
    float rsum = 0.f, dval = 0.f;
    bool dfound = false;
    float* colpart = args.colpart + (int64_t)blockIdx.x * args.Nb;
    const uint32_t e_base = smem_u32(e_tiles);
    auto flush_cols = [&](int t) {   // column sums of tile t: TMEM lane == column; read by four warps of row half 0
      const int buf = t & 1;
      mbar_wait(&c_full[buf], (uint32_t)(t >> 1) & 1u);
      tc_fence_after_sync();
      uint32_t v;
      asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=r"(v)
                   : "r"(tmem_c + ((uint32_t)(q * 32) << 16) + (uint32_t)buf * 16u) : "memory");
      tmem_ld_wait();
      const int64_t col = (int64_t)(tile_begin + t) * kTile + i;
      if (col < args.Nb) colpart[col] = __uint_as_float(v);
      tc_fence_before_sync();
      mbar_arrive(&c_empty[buf]);
    };
    if (grp < nh) {
      const int h = grp;
      const bool flusher = grp == 0 && colhalf == 0;
      const int64_t rb = row0 + h * kTile;                          // first row of this half
      const int64_t row = rb + i;
      const int64_t drow = args.diag_offset + row;                  // my positive's column
      const int64_t dlo = args.diag_offset + rb, dhi = dlo + kTile; // diagonal band of this half
      const bool row_ragged = rb + kTile > args.Ma;
      for (int t = 0; t < ntiles; ++t) {
        const int u = t * nh + h;
        const int sb = u % kSBuf, eb = u % kEBuf;
        const uint32_t sph = (uint32_t)(u / kSBuf) & 1u, eph = (uint32_t)(u / kEBuf) & 1u;
        const int64_t col0 = (int64_t)(tile_begin + t) * kTile;
        const bool special = row_ragged || (col0 + kTile > args.Nb) || (col0 < dhi && col0 + kTile > dlo);
        mbar_wait(&e_empty[eb], eph ^ 1u);   // the column-sum MMA of item u - kEBuf has consumed this buffer
        mbar_wait(&s_full[sb], sph);
        tc_fence_after_sync();
        if (threadIdx.x == 64) PCLIP_TRACE(3, u);
        const uint32_t panel = e_base + (uint32_t)eb * kTileBytes + (uint32_t)colhalf * kPanelBytes;
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          uint32_t v[32];
          tmem_ld_32x32b_x32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)sb * kTile +
                                 (uint32_t)(colhalf * 64 + c * 32), v);
          tmem_ld_wait();
          if (c == 1) {
            tc_fence_before_sync();
            mbar_arrive(&s_empty[sb]);              // S(u) is in registers: the MMA warp may overwrite the accumulator
            if (threadIdx.x == 64) PCLIP_TRACE(4, u);
          }
          uint32_t packed[16];
          if (!special) {
            uint64_t acc = pack2(0.f, 0.f);
#pragma unroll
            for (int e = 0; e < 32; e += 2) {
              float e0, e1;
              if (e < kPoly) {
                exp_poly2<4>(__uint_as_float(v[e]), __uint_as_float(v[e + 1]), e0, e1);
              } else {
                float y0, y1;
                unpack2(mul2(pack2(__uint_as_float(v[e]), __uint_as_float(v[e + 1])), pack2(kLog2e, kLog2e)), y0, y1);
                e0 = ex2_approx(y0);
                e1 = ex2_approx(y1);
              }
              acc = add2(acc, pack2(e0, e1));
              packed[e >> 1] = bf16x2_bits(e0, e1);
            }
            float a0, a1;
            unpack2(acc, a0, a1);
            rsum += a0 + a1;
          } else {
#pragma unroll
            for (int e = 0; e < 32; e += 2) {
              float ev[2];
#pragma unroll
              for (int k2 = 0; k2 < 2; ++k2) {
                const int64_t col = col0 + colhalf * 64 + c * 32 + e + k2;
                const float sv = __uint_as_float(v[e + k2]);
                float x = (col < args.Nb && row < args.Ma) ? ex2_approx(sv * kLog2e) : 0.f;
                if (col < args.Nb) rsum += x;
                if (col == drow) {       // keep the positive out of the bf16 column sum: col_reduce adds it in fp32
                  dval = sv;
                  dfound = true;
                  x = 0.f;
                }
                ev[k2] = x;
              }
              packed[e >> 1] = bf16x2_bits(ev[0], ev[1]);
            }
          }
#pragma unroll
          for (int g = 0; g < 4; ++g)
            st_shared_v4(panel + sw128_offset((uint32_t)i, (uint32_t)(c * 4 + g)), packed[4 * g], packed[4 * g + 1],
                         packed[4 * g + 2], packed[4 * g + 3]);
        }
        if (threadIdx.x == 64) PCLIP_TRACE(5, u);
        fence_proxy_async_smem();
        mbar_arrive(&e_full[eb]);
        if (threadIdx.x == 64) PCLIP_TRACE(6, u);
        if (flusher && t > 0) flush_cols(t - 1);   // one tile behind: never waits on the MMA just requested
      }
      if (flusher && ntiles > 0) flush_cols(ntiles - 1);
      // row sums: one partial per (split, column quarter); this thread covered two quarters
      if (row < args.Ma) {
        args.rowpart[((int64_t)blockIdx.y * 4 + colhalf * 2) * args.Ma + row] = rsum;
        args.rowpart[((int64_t)blockIdx.y * 4 + colhalf * 2 + 1) * args.Ma + row] = 0.f;
        if (dfound) args.diag[row] = dval;
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

// =================================================================================================================
// host side
// =================================================================================================================
__global__ void grad_reduce_kernel(GradReduceSide s0, GradReduceSide s1, int nsplit, int64_t Ma, float scale,
                                   const float* __restrict__ gscale);
__global__ void sum_partials_kernel(const float* __restrict__ partial, int nsplit, int64_t M, float* __restrict__ out);

// colsum[j] = sum_r colpart[r][j] + exp(diag[j - off]) for the columns that are some local row's positive
__global__ void col_reduce_kernel(const float* __restrict__ colpart, int nrb, int64_t Nb, const float* __restrict__ diag,
                                  int64_t off, int64_t Ma, float* __restrict__ colsum) {
  const int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (j >= Nb) return;
  float s = 0.f;
  int r = 0;
  for (; r + 8 <= nrb; r += 8) {           // eight loads in flight, added in row-block order
    float v[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) v[u] = __ldcs(colpart + (int64_t)(r + u) * Nb + j);
#pragma unroll
    for (int u = 0; u < 8; ++u) s += v[u];
  }
  for (; r < nrb; ++r) s += colpart[(int64_t)r * Nb + j];
  if (j >= off && j < off + Ma) s += expf(diag[j - off]);
  colsum[j] = s;
}

// Column splits: the CTA count should fill whole waves of the SMs (one CTA per SM) without paying the per-CTA fixed cost
// too often.  Model: time ~ waves(ns) * (tiles per split + overhead), overhead = the prologue / final epilogue / partial
// reduction of one CTA in units of one column tile (measured on the rank-local problem of an 8-GPU run,
// tools/gpu_shard_sweep.py: ~8 tiles for the gradient kernel, ~2 for the forward kernel).  512 row blocks on 148 SMs:
// 2 splits (6.9 waves of 264 instead of 3.5 waves -> 4 of 520).  A B/8 row shard (64 CTAs): 9 splits (3.9 waves of 65),
// not "as many as fill the last wave best" (23 splits of 23 tiles measured 20 % slower).
static void plan_waves(int64_t row_blocks, int64_t ntiles, int n_sides, int max_split, int overhead_tiles, int* nsplit,
                       int* tps) {
  const int sms = std::max(device_sm_count(), 1);
  const int64_t base = row_blocks * n_sides;
  const int cap = base < sms ? 8 * max_split : max_split;
  int64_t best = -1;
  int best_ns = 1;
  for (int ns = 1; ns <= cap && (ns == 1 || ns * 4 <= ntiles); ++ns) {
    const int64_t t = ceil_div(ntiles, (int64_t)ns);                       // tiles per split
    const int64_t ctas = base * ceil_div(ntiles, t);
    const int64_t cost = ceil_div(ctas, (int64_t)sms) * (t + overhead_tiles);
    if (best < 0 || cost < best) {
      best = cost;
      best_ns = ns;
    }
  }
  if (const char* e = getenv("PCLIP_NSPLIT")) {   // tuning override (tools/gpu_shard_sweep.py)
    const int v = atoi(e);
    if (v >= 1 && v <= cap && (v == 1 || v * 4 <= ntiles)) best_ns = v;
  }
  *tps = (int)ceil_div(ntiles, (int64_t)best_ns);
  *nsplit = (int)ceil_div(ntiles, (int64_t)*tps);
}

size_t tc2_grad_workspace(int64_t Ma, int64_t Nb) {
  int ns, ns2, tps;
  plan_waves(ceil_div(Ma, c2::kRows), ceil_div(Nb, c2::kTile), 1, 4, 8, &ns, &tps);
  plan_waves(ceil_div(Ma, c2::kRows), ceil_div(Nb, c2::kTile), 2, 4, 8, &ns2, &tps);
  ns = std::max(ns, ns2);
  return ns > 1 ? 2 * align_up((size_t)ns * Ma * c2::kDim * sizeof(float), 1024) + 2048 : 2048;
}

int tc2_grad(const pclip_clip_bwd_side_t* sides, int n_sides, int D, float scale, const float* gscale, void* ws,
             size_t ws_bytes, cudaStream_t st) {
  using namespace c2;
  PCLIP_REQUIRE(D == kDim, PCLIP_ERR_UNSUPPORTED, "bf16 tensor-core path needs D == %d (got %d)", kDim, D);
  for (int i = 0; i < n_sides; ++i) {
    const pclip_clip_bwd_side_t& s = sides[i];
    PCLIP_REQUIRE(s.a && s.b && s.rinv && s.cinv && s.da && s.Ma > 0 && s.Nb > 0, PCLIP_ERR_INVALID, "grad: bad side %d", i);
    PCLIP_REQUIRE(s.diag_offset >= 0 && s.diag_offset + s.Ma <= s.Nb, PCLIP_ERR_INVALID,
                  "grad: positives of side %d fall outside b", i);
    PCLIP_REQUIRE(s.Ma == sides[0].Ma && s.Nb == sides[0].Nb, PCLIP_ERR_UNSUPPORTED,
                  "grad: both sides must have the same shape");
  }
  const int64_t Ma = sides[0].Ma, Nb = sides[0].Nb;
  int nsplit, tps;
  plan_waves(ceil_div(Ma, kRows), ceil_div(Nb, kTile), n_sides, 4, 8, &nsplit, &tps);
  Workspace w(ws, ws_bytes);
  Tc2BwdArgs args;
  CUtensorMap maps[4];
  float* partial[2] = {nullptr, nullptr};
  for (int i = 0; i < 2; ++i) {
    const pclip_clip_bwd_side_t& s = sides[i < n_sides ? i : 0];
    PCLIP_TRY(make_tmap_bf16_2d(&maps[2 * i], s.a, (uint64_t)s.Ma, kDim, kDim, kTile, 64));
    PCLIP_TRY(make_tmap_bf16_2d(&maps[2 * i + 1], s.b, (uint64_t)s.Nb, kDim, kDim, kTile, 64));
    if (nsplit > 1 && i < n_sides) {
      PCLIP_REQUIRE(ws, PCLIP_ERR_WORKSPACE, "grad: workspace required");
      partial[i] = w.take<float>((size_t)nsplit * s.Ma * kDim);
    }
    args.s[i] = {s.Ma,   s.Nb,   s.diag_offset, s.rinv, s.cinv, static_cast<const __nv_bfloat16*>(s.b),
                 nsplit > 1 ? partial[i] : s.da};
  }
  PCLIP_REQUIRE(nsplit == 1 || w.ok(), PCLIP_ERR_WORKSPACE, "grad: workspace too small");
  args.tiles_per_split = tps;
  args.direct = nsplit == 1;
  args.scale = scale;
  args.gscale = gscale;
  static const char* trace_path = getenv("PCLIP_BWD_TRACE");
  args.trace = nullptr;
  if (trace_path) {
    PCLIP_CHECK_CUDA(cudaMalloc(&args.trace, 8 * kTraceItems * sizeof(long long)));
    PCLIP_CHECK_CUDA(cudaMemsetAsync(args.trace, 0, 8 * kTraceItems * sizeof(long long), st));
  }
  dim3 grid((unsigned)ceil_div(Ma, kRows), (unsigned)nsplit, (unsigned)n_sides);
  PCLIP_CHECK_CUDA(cudaFuncSetAttribute(tc3_grad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)c3::kSmem3));
  tc3_grad_kernel<<<grid, c3::kThreads3, c3::kSmem3, st>>>(maps[0], maps[1], maps[2], maps[3], args);
  PCLIP_CHECK_LAUNCH();
  if (args.trace) {   // debug only: synchronises
    static long long h[8 * kTraceItems];
    PCLIP_CHECK_CUDA(cudaStreamSynchronize(st));
    PCLIP_CHECK_CUDA(cudaMemcpy(h, args.trace, sizeof(h), cudaMemcpyDeviceToHost));
    cudaFree(args.trace);
    if (FILE* f = fopen(trace_path, "w")) {
      const long long t0 = h[0];
      fprintf(f, "# item(2t+h) s_issue pv_issue epi_start epi_sfull epi_chunk0_math pvdone_passed epi_done (clocks since S(0,0))\n");
      for (int u = 0; u < kTraceItems; ++u) {
        fprintf(f, "%3d", u);
        for (int ev = 0; ev < 7; ++ev) fprintf(f, " %8lld", h[ev * kTraceItems + u] ? h[ev * kTraceItems + u] - t0 : -1);
        fprintf(f, "\n");
      }
      fclose(f);
    }
  }
  if (nsplit > 1) {
    GradReduceSide r[2];
    for (int i = 0; i < 2; ++i) {
      const pclip_clip_bwd_side_t& sd = sides[i < n_sides ? i : 0];
      r[i] = {partial[i < n_sides ? i : 0], sd.b, sd.da, sd.diag_offset};
    }
    grad_reduce_kernel<<<dim3((unsigned)ceil_div(Ma * (kDim / 4), 256), (unsigned)n_sides), 256, 0, st>>>(
        r[0], r[1], nsplit, Ma, scale, gscale);
    PCLIP_CHECK_LAUNCH();
  }
  return PCLIP_OK;
}

static void plan_fwd2(int64_t Ma, int64_t Nb, int* nsplit, int* tps) {
  plan_waves(ceil_div(Ma, c2::kRows), ceil_div(Nb, c2::kTile), 1, 8, 2, nsplit, tps);
}

size_t tc2_sumexp_workspace(int64_t Ma, int64_t Nb) {
  int ns, tps;
  plan_fwd2(Ma, Nb, &ns, &tps);
  return align_up((size_t)4 * ns * Ma * sizeof(float), 1024) +
         align_up((size_t)ceil_div(Ma, c2::kRows) * Nb * sizeof(float), 1024) + 2048;
}

int tc2_sumexp(const void* a, const void* b, int64_t Ma, int64_t Nb, int64_t diag_offset, int D, float* rowsum,
               float* colsum, float* diag, void* ws, size_t ws_bytes, cudaStream_t st) {
  using namespace c2;
  PCLIP_REQUIRE(D == kDim, PCLIP_ERR_UNSUPPORTED, "bf16 tensor-core path needs D == %d (got %d)", kDim, D);
  int nsplit, tps;
  plan_fwd2(Ma, Nb, &nsplit, &tps);
  const int nrb = (int)ceil_div(Ma, kRows);
  Workspace w(ws, ws_bytes);
  float* rowpart = w.take<float>((size_t)4 * nsplit * Ma);
  float* colpart = w.take<float>((size_t)nrb * Nb);
  PCLIP_REQUIRE(ws && w.ok(), PCLIP_ERR_WORKSPACE, "sumexp: workspace too small");
  CUtensorMap ma, mb;
  PCLIP_TRY(make_tmap_bf16_2d(&ma, a, (uint64_t)Ma, kDim, kDim, kTile, 64));
  PCLIP_TRY(make_tmap_bf16_2d(&mb, b, (uint64_t)Nb, kDim, kDim, kTile, 64));
  static const char* trace_path = getenv("PCLIP_FWD_TRACE");
  long long* trace = nullptr;
  if (trace_path) {
    PCLIP_CHECK_CUDA(cudaMalloc(&trace, 8 * kTraceItems * sizeof(long long)));
    PCLIP_CHECK_CUDA(cudaMemsetAsync(trace, 0, 8 * kTraceItems * sizeof(long long), st));
  }
  Tc2FwdArgs args{Ma, Nb, diag_offset, rowpart, colpart, diag, tps, trace};
  dim3 grid((unsigned)nrb, (unsigned)nsplit);
  // exps per 32 on the FMA pipes (tuning knob; the default was chosen from B200 measurements, profiles/)
  static const int poly = [] {
    const char* e = getenv("PCLIP_FWD_POLY");
    return e ? atoi(e) : kFwdPoly;
  }();
#define PCLIP_LAUNCH_FWD(P)                                                                                          \
  do {                                                                                                               \
    PCLIP_CHECK_CUDA(cudaFuncSetAttribute(tc2_sumexp_kernel<P>, cudaFuncAttributeMaxDynamicSharedMemorySize,         \
                                          (int)kFwdSmem));                                                           \
    tc2_sumexp_kernel<P><<<grid, kFwdThreads, kFwdSmem, st>>>(ma, mb, args);                                         \
  } while (0)
  if (poly <= 0) PCLIP_LAUNCH_FWD(0);
  else if (poly <= 6) PCLIP_LAUNCH_FWD(6);
  else if (poly <= 8) PCLIP_LAUNCH_FWD(8);
  else if (poly <= 10) PCLIP_LAUNCH_FWD(10);
  else if (poly <= 12) PCLIP_LAUNCH_FWD(12);
  else if (poly <= 14) PCLIP_LAUNCH_FWD(14);
  else PCLIP_LAUNCH_FWD(18);
#undef PCLIP_LAUNCH_FWD
  PCLIP_CHECK_LAUNCH();
  if (trace) {   // debug only: synchronises
    static long long h[8 * kTraceItems];
    PCLIP_CHECK_CUDA(cudaStreamSynchronize(st));
    PCLIP_CHECK_CUDA(cudaMemcpy(h, trace, sizeof(h), cudaMemcpyDeviceToHost));
    cudaFree(trace);
    if (FILE* f = fopen(trace_path, "w")) {
      long long t0 = h[1 * kTraceItems];
      fprintf(f, "# item c_issued s_issue c_issue epi_sfull epi_loaded epi_math_done epi_done s_issued (clocks since S(0) issue)\n");
      for (int u = 0; u < kTraceItems; ++u) {
        fprintf(f, "%3d", u);
        for (int ev : {0, 1, 2, 3, 4, 5, 6, 7}) fprintf(f, " %8lld", h[ev * kTraceItems + u] ? h[ev * kTraceItems + u] - t0 : -1);
        fprintf(f, "\n");
      }
      fclose(f);
    }
  }
  sum_partials_kernel<<<(unsigned)ceil_div(Ma, 256), 256, 0, st>>>(rowpart, 4 * nsplit, Ma, rowsum);
  PCLIP_CHECK_LAUNCH();
  col_reduce_kernel<<<(unsigned)ceil_div(Nb, 256), 256, 0, st>>>(colpart, nrb, Nb, diag, diag_offset, Ma, colsum);
  PCLIP_CHECK_LAUNCH();
  return PCLIP_OK;
}

}  // namespace pclip
