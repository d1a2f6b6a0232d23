// FILIP forward in ONE sweep over the score tiles (PCLIP_COMPUTE_BF16, D == 128, row side with at most 512 tokens).
//   ref: modules/models.py:332-351 (amax over the column tokens AND amax over the row tokens of the same S)
// filip_tc.cu computes S once per side (S for the row maxima of A, S^T for the row maxima of B: thread == TMEM lane ==
// token row, so only the row direction is in-thread).  Here every 128 x 128 score tile is produced once and read with
// tcgen05.ld.16x256b instead of .32x32b: a thread then holds a 4-row x 8-column patch of each 32 x 32 chunk (rows
// g + 8j, columns 8k + 2c + e of the warp's 32 TMEM lanes; g = lane / 4, c = lane % 4 -- the mma.m16n8 accumulator
// layout), so BOTH directions start with an in-thread reduction:
//   row direction     8 keys of a row -> 1 in-thread (FMNMX3 tree); the chunk winners of a tile are merged by plain
//                     maxima (the chunk id rides in the key's j bits) and compared with the running maximum per
//                     (row, lane) once per tile; the 4 lanes that share a row are combined once per sample pair
//   column direction  4 keys of a column -> 1 in-thread; the 8 lane groups g that share a column set exchange their
//                     candidates through a warp-private 1 KB shared-memory patch (conflict-free [column slot][g]
//                     layout), which leaves one column winner per lane; the winners of the 16 (row tile, lane quarter)
//                     warps meet in shared memory, where a finaliser warp picks the maximiser over all rows of the
//                     sample (FMNMX3 tree over slot-tagged keys) and writes argY / the masked-mean partial sum
// One key serves both directions: the score with its 5 low mantissa bits replaced by 31 - (8j + 2k + e) (ONE LOP3);
// within a row j is constant, within a column (k, e) is.  Before the exchange the 3 column bits are replaced by
// 7 - g, so the 5 bits of a column winner are 31 - (row within the warp's 32 lanes).  Tie rule as in filip_tc.cu:
// scores equal to 2^-18 relative are tied and the lowest index wins.
// The CTA owns ALL rows of one row sample (up to 4 row tiles stationary in shared memory; the 4 TMEM accumulators are
// a ring over the (column tile, row tile) sequence), so the column maxima are complete inside the CTA.  Warps: TMA
// producer, MMA issuer, 16 epilogue warps (row tile h x lane quarter q), column finaliser.  The kernel is bound by the
// ALU pipe (LOP3 / FMNMX: 81 % active), then by the power-limited tensor rate; DESIGN.md section 4 has the measured history.
#include <cuda_bf16.h>
#include <float.h>

#include "common.cuh"
#include "ptx.cuh"

namespace pclip {

namespace f1 {
constexpr int kTile = 128;
constexpr int kDim = 128;
constexpr uint32_t kPanelBytes = 128 * 128;
constexpr uint32_t kTileBytes = 2 * kPanelBytes;
constexpr int kMaxH = 4;                       // row tiles per CTA
constexpr int kStages = 2;                     // column-tile ring (each tile feeds up to 4 MMAs)
constexpr int kThreads = 32 * (2 + 4 * kMaxH + 1);   // TMA warp, MMA warp, 4 epilogue warps per row tile, column finaliser
constexpr int kSlots = kMaxH * 4;              // (row tile, lane quarter) writers per column tile
constexpr uint32_t kColBufBytes = 2 * kSlots * kTile * 4;
constexpr uint32_t kScratchBytes = 4 * kMaxH * 1024;   // one 32 x 8-word transposition patch per epilogue warp
constexpr uint32_t kSmem = kTileBytes * (kMaxH + kStages) + kColBufBytes + kScratchBytes + 256 + 1024;
static_assert(kSmem <= 232448, "shared memory");
constexpr uint32_t kNegKey = 0xff7fffe0u;      // key of -FLT_MAX
constexpr uint32_t kNegMax = 0xff7fffffu;      // -FLT_MAX

__device__ __forceinline__ void issue_tile(uint32_t d, uint32_t a_addr, uint32_t b_addr, uint32_t idesc) {
  const uint64_t ad = make_smem_desc_sw128(a_addr, 16, 1024), bd = make_smem_desc_sw128(b_addr, 16, 1024);
  if (elect_one()) {
#pragma unroll
    for (int k = 0; k < kDim / 16; ++k) {
      const uint32_t off = (uint32_t)(k >> 2) * kPanelBytes + (uint32_t)(k & 3) * 32u;
      umma_bf16_ss(d, desc_advance(ad, off), desc_advance(bd, off), idesc, k > 0 ? 1u : 0u);
    }
  }
  __syncwarp();
}

// 16 TMEM lanes x 32 columns: r[4k + 2jj + e] = (lane base + g + 8 jj, column 8k + 2c + e)
__device__ __forceinline__ void tmem_ld_16x256b_x4(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.16x256b.x4.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}

// a 32-row x 32-column chunk of the warp's lane quarter:
//   r[16 (j >> 1) + 4k + 2 (j & 1) + e] = (row g + 8j, column 8k + 2c + e)
__device__ __forceinline__ void load_chunk(uint32_t taddr, uint32_t (&r)[32]) {
  tmem_ld_16x256b_x4(taddr, &r[0]);
  tmem_ld_16x256b_x4(taddr + (16u << 16), &r[16]);
}

__device__ __forceinline__ constexpr int reg_of(int j, int k, int e) { return 16 * (j >> 1) + 4 * k + 2 * (j & 1) + e; }

template <int I>
__device__ __forceinline__ void make_keys(const uint32_t (&r)[32], const uint32_t keep, float (&key)[32]) {
  if constexpr (I < 32) {
    constexpr int j = 2 * (I >> 4) + ((I >> 1) & 1), k = (I >> 2) & 3, e = I & 1;
    uint32_t v;
    asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(v) : "r"(r[I]), "r"(keep), "n"(31 - (8 * j + 2 * k + e)));   // (r & keep) | code
    key[I] = __uint_as_float(v);
    make_keys<I + 1>(r, keep, key);
  }
}

__device__ __forceinline__ float max3(float a, float b, float c) { return fmaxf(fmaxf(a, b), c); }   // one FMNMX3

struct RowState {
  uint32_t mk[4];   // running maximum of row g + 8j over this lane's columns, as a key
  int base[4];      // first column of the tile it came from
};

__device__ __forceinline__ void reset(RowState& s) {
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    s.mk[j] = kNegKey;
    s.base[j] = 0;
  }
}

// After a tile's chunks: the tile's row winners against the running maxima (an earlier tile keeps a tie).
__device__ __forceinline__ void tile_rows(const float (&rowwin)[4], const int s0, RowState& st) {
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const uint32_t m = __float_as_uint(rowwin[j]);
    if (__uint_as_float(m | 31u) > __uint_as_float(st.mk[j] | 31u)) {
      st.mk[j] = m;
      st.base[j] = s0;
    }
  }
}

struct Patch {         // this lane's addresses in its warp's transposition patch
  uint32_t* wr;        // word 8c + g
  const uint4* rd0;    // slot `lane`, the half lanes with bit 2 clear read first
  const uint4* rd1;
  uint32_t gcode;      // 7 - g
};

// One 32 x 32 chunk (chunk cc of its tile).  nv = valid columns of the chunk (>= 1; < 32 only in a column sample's
// last tile), rowmask: bit j set = row g + 8j is a pad row (excluded from the column maxima).  Returns the winner of column
// 8 (lane >> 3) + 2 (lane & 3) + ((lane >> 2) & 1) of the chunk, low 5 bits = 31 - (winning row within the warp's 32).
__device__ __forceinline__ uint32_t chunk_both(uint32_t (&r)[32], const int cc, const int nv, const uint32_t rowmask,
                                               const int lane, float (&rowwin)[4], const Patch& pa) {
  const int c = lane & 3;
  if (nv < 32) {
#pragma unroll
    for (int k = 0; k < 4; ++k)
#pragma unroll
      for (int e = 0; e < 2; ++e)
        if (8 * k + 2 * c + e >= nv) {
#pragma unroll
          for (int j = 0; j < 4; ++j) r[reg_of(j, k, e)] = kNegMax;
        }
  }
  float key[32];
  uint32_t keep;   // in a register the optimiser cannot see through: (r & keep) | imm is then ONE LOP3
  asm("mov.u32 %0, 0xffffffe0;" : "=r"(keep));
  make_keys<0>(r, keep, key);
  // row direction: 8 keys -> 1; the j bits (constant within a row) are replaced by 3 - (chunk within the tile), so the
  // chunk winners of a tile combine by plain maxima and the running maximum is updated once per tile (tile_rows)
  uint32_t cmask, ccode = (uint32_t)(3 - cc) << 3;
  asm("mov.u32 %0, 0xffffffe7;" : "=r"(cmask));
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float a = max3(key[reg_of(j, 0, 0)], key[reg_of(j, 0, 1)], key[reg_of(j, 1, 0)]);
    const float b = max3(key[reg_of(j, 1, 1)], key[reg_of(j, 2, 0)], key[reg_of(j, 2, 1)]);
    const uint32_t m = __float_as_uint(max3(a, b, fmaxf(key[reg_of(j, 3, 0)], key[reg_of(j, 3, 1)])));
    uint32_t v;
    asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(v) : "r"(m), "r"(cmask), "r"(ccode));
    rowwin[j] = cc == 0 ? __uint_as_float(v) : fmaxf(rowwin[j], __uint_as_float(v));
  }
  // column direction: pad rows out first
  if (rowmask) {
#pragma unroll
    for (int j = 0; j < 4; ++j)
      if (rowmask & (1u << j)) {
#pragma unroll
        for (int k = 0; k < 4; ++k)
#pragma unroll
          for (int e = 0; e < 2; ++e) key[reg_of(j, k, e)] = __uint_as_float(kNegKey);
      }
  }
  // 4 rows -> 1 in-thread, tagged with 7 - g in the 3 bits the column index no longer needs; then the 8 lanes (g) that
  // share a column set meet through a warp-private shared-memory patch [column slot][g] (1 KB): slot = c | e << 2 |
  // k << 3, so the 32 lanes of one store hit 32 banks (word 8c + g + 32e + 64k), and lane L reads slot L's 8
  // candidates with two 16-byte loads (lanes with bit 2 set take the halves in the other order: no bank conflicts)
  // and reduces them with one FMNMX3 tree.  (The first version exchanged the candidates with a transposing shuffle
  // butterfly: 14 SEL + 7 FMNMX more on the ALU pipe that bounds this kernel.)
  uint32_t gkeep;
  asm("mov.u32 %0, 0xfffffff8;" : "=r"(gkeep));
#pragma unroll
  for (int k = 0; k < 4; ++k)
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const float m = max3(key[reg_of(0, k, e)], key[reg_of(1, k, e)], fmaxf(key[reg_of(2, k, e)], key[reg_of(3, k, e)]));
      uint32_t v;
      asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(v) : "r"(__float_as_uint(m)), "r"(gkeep), "r"(pa.gcode));
      pa.wr[32 * e + 64 * k] = v;
    }
  __syncwarp();
  const uint4 a = *pa.rd0, b = *pa.rd1;
  __syncwarp();
  return __float_as_uint(max3(max3(__uint_as_float(a.x), __uint_as_float(a.y), __uint_as_float(a.z)),
                              max3(__uint_as_float(a.w), __uint_as_float(b.x), __uint_as_float(b.y)),
                              fmaxf(__uint_as_float(b.z), __uint_as_float(b.w))));
}
}  // namespace f1

struct FilipBothArgs {
  int Bx, By, Tx, Ty;
  const int32_t* lenX;
  const int32_t* lenY;
  int y_per_split;
  float* partX;   // [Bx][By][npX]: slot = row tile * 4 + lane quarter
  int npX;
  int16_t* argX;  // [Bx][By][Tx]
  float* partY;   // [By][Bx]
  int16_t* argY;  // [By][Bx][Ty]
};

// grid (Bx, ysplit).  kDbg (PCLIP_FILIP_DBG, timing experiments only): 1 = no epilogue arithmetic, 2 = no TMEM loads either
template <int kDbg>
__global__ void __launch_bounds__(f1::kThreads, 1)
tc_filip_both_kernel(const __grid_constant__ CUtensorMap tx, const __grid_constant__ CUtensorMap ty,
                     const FilipBothArgs args) {
  using namespace f1;
  const int bx = blockIdx.x;
  const int y0 = blockIdx.y * args.y_per_split, y1 = min(y0 + args.y_per_split, args.By);
  const int lx = max(min(args.lenX[bx], args.Tx), 0);
  const int nh = (lx + kTile - 1) / kTile;   // row tiles that hold at least one valid token (0..4)

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* a_tiles = smem;
  uint8_t* b_tiles = a_tiles + kMaxH * kTileBytes;
  uint32_t* colbuf = reinterpret_cast<uint32_t*>(b_tiles + kStages * kTileBytes);   // [2][kSlots][128]
  uint32_t* scratch_all = colbuf + kColBufBytes / 4;                                 // [16 warps][32 slots][8]
  uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(scratch_all) + kScratchBytes);
  uint64_t* a_full = bars;
  uint64_t* b_full = a_full + 1;            // [kStages]
  uint64_t* b_empty = b_full + kStages;     // [kStages]
  uint64_t* t_full = b_empty + kStages;     // [4]
  uint64_t* t_empty = t_full + 4;           // [4]
  uint64_t* c_full = t_empty + 4;           // [2]
  uint64_t* c_empty = c_full + 2;           // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(c_empty + 2);

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    mbar_init(a_full, 1);
    for (int i = 0; i < kStages; ++i) {
      mbar_init(&b_full[i], 1);
      mbar_init(&b_empty[i], 1);
    }
    for (int i = 0; i < 4; ++i) {
      mbar_init(&t_full[i], 1);
      mbar_init(&t_empty[i], 128);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&c_full[i], (uint32_t)max(nh, 1) * 128u);
      mbar_init(&c_empty[i], 32);
    }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc<512>(tmem_slot);
  // slots of row tiles without a valid token are never written: they must lose every comparison
  for (int i = nh * 4 * kTile + (int)threadIdx.x; i < kSlots * kTile; i += kThreads) {
    colbuf[i] = kNegKey;
    colbuf[kSlots * kTile + i] = kNegKey;
  }
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
        tma_load_3d(a_tiles + h * kTileBytes, &tx, a_full, 0, h * kTile, bx);
        tma_load_3d(a_tiles + h * kTileBytes + kPanelBytes, &tx, a_full, 64, h * kTile, bx);
      }
      int ct = 0;
      for (int by = y0; by < y1; ++by) {
        const int ly = min(args.lenY[by], args.Ty);
        for (int s0 = 0; s0 < ly; s0 += kTile, ++ct) {
          const int stage = ct % kStages;
          const uint32_t phase = (uint32_t)(ct / kStages) & 1u;
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
      int ct = 0, it = 0;
      for (int by = y0; by < y1; ++by) {
        const int ly = min(args.lenY[by], args.Ty);
        for (int s0 = 0; s0 < ly; s0 += kTile, ++ct) {
          const int stage = ct % kStages;
          const uint32_t phase = (uint32_t)(ct / kStages) & 1u;
          mbar_wait(&b_full[stage], phase);
          const uint32_t b_addr = smem_u32(b_tiles + (size_t)stage * kTileBytes);
          for (int h = 0; h < nh; ++h, ++it) {
            const int buf = it & 3;
            mbar_wait(&t_empty[buf], ((uint32_t)(it >> 2) & 1u) ^ 1u);
            tc_fence_after_sync();
            issue_tile(tmem_base + (uint32_t)buf * kTile, smem_u32(a_tiles + h * kTileBytes), b_addr, idesc);
            if (elect_one()) {
              umma_commit(&t_full[buf]);
              if (h == nh - 1) umma_commit(&b_empty[stage]);
            }
            __syncwarp();
          }
        }
      }
    }
  } else if (warp < 2 + 4 * kMaxH) {
    // ---- epilogue: warp (h, q) reads lane quarter q of row tile h ----
    const int q = warp & 3, h = (warp - 2) >> 2;
    const int g = lane >> 2, c = lane & 3;
    const bool live = h < nh;
    uint32_t* scratch = scratch_all + (warp - 2) * 256;
    const int hb = (lane >> 2) & 1;
    const Patch pa{scratch + 8 * c + g, reinterpret_cast<const uint4*>(scratch + 8 * lane + 4 * hb),
                   reinterpret_cast<const uint4*>(scratch + 8 * lane + 4 * (1 - hb)), 7u - (uint32_t)g};
    RowState st;
    reset(st);
    uint32_t rowmask = 0u;
#pragma unroll
    for (int j = 0; j < 4; ++j)
      if (h * kTile + q * 32 + g + 8 * j >= lx) rowmask |= 1u << j;
    int ct = 0;
    for (int by = y0; by < y1; ++by) {
      const int ly = min(args.lenY[by], args.Ty);
      if (live) {
        for (int s0 = 0; s0 < ly; s0 += kTile, ++ct) {
          const int it = ct * nh + h;
          const int buf = it & 3;
          mbar_wait(&t_full[buf], (uint32_t)(it >> 2) & 1u);
          tc_fence_after_sync();
          const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)buf * kTile;
          const int nvalid = min(ly - s0, kTile);   // >= 1
          const int nb = ct & 1;
          uint32_t* mine = colbuf + ((size_t)nb * kSlots + h * 4 + q) * kTile + 8 * (lane >> 3) + 2 * c + ((lane >> 2) & 1);
          uint32_t ra[32];
          float rowwin[4];
          if (nvalid == kTile) {
            uint32_t w[4];
#pragma unroll
            for (int cc = 0; cc < 4; ++cc) {
              if (kDbg < 2) load_chunk(taddr + cc * 32, ra);
              tmem_ld_wait();
              if (cc == 3) {   // the accumulator goes back to the MMA warp as soon as the last load has landed
                tc_fence_before_sync();
                mbar_arrive(&t_empty[buf]);
              }
              w[cc] = kDbg ? ra[cc] : chunk_both(ra, cc, 32, rowmask, lane, rowwin, pa);
            }
            if (!kDbg) tile_rows(rowwin, s0, st);
            mbar_wait(&c_empty[nb], ((uint32_t)(ct >> 1) & 1u) ^ 1u);
#pragma unroll
            for (int cc = 0; cc < 4; ++cc) mine[cc * 32] = w[cc];
          } else {
            mbar_wait(&c_empty[nb], ((uint32_t)(ct >> 1) & 1u) ^ 1u);
#pragma unroll 1
            for (int cc = 0; cc * 32 < nvalid; ++cc) {   // warp-uniform
              load_chunk(taddr + cc * 32, ra);
              tmem_ld_wait();
              mine[cc * 32] = chunk_both(ra, cc, min(nvalid - cc * 32, 32), rowmask, lane, rowwin, pa);
            }
            tile_rows(rowwin, s0, st);
            tc_fence_before_sync();
            mbar_arrive(&t_empty[buf]);
          }
          mbar_arrive(&c_full[nb]);
        }
      }
      // ---- end of the sample pair: the 4 lanes that share a row (highest score bits, then lowest column) ----
      uint32_t sb[4];
      int col[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int v = 7 - (int)(st.mk[j] & 7u), cc = 3 - (int)((st.mk[j] >> 3) & 3u);
        sb[j] = st.mk[j] | 31u;
        col[j] = st.base[j] + 32 * cc + 8 * (v >> 1) + 2 * c + (v & 1);
#pragma unroll
        for (int o = 1; o <= 2; o <<= 1) {
          const uint32_t osb = __shfl_xor_sync(0xffffffffu, sb[j], o);
          const int ocol = __shfl_xor_sync(0xffffffffu, col[j], o);
          const bool take = __uint_as_float(osb) > __uint_as_float(sb[j]) || (osb == sb[j] && ocol < col[j]);
          sb[j] = take ? osb : sb[j];
          col[j] = take ? ocol : col[j];
        }
      }
      const uint32_t msb = c == 0 ? sb[0] : c == 1 ? sb[1] : c == 2 ? sb[2] : sb[3];
      const int mcol = c == 0 ? col[0] : c == 1 ? col[1] : c == 2 ? col[2] : col[3];
      const int t = h * kTile + q * 32 + g + 8 * c;   // this lane reports row g + 8c
      // the 5 bits the index displaced are set to the middle of their range: |error| <= 2^-19 relative, unbiased
      const float m = ly > 0 ? __uint_as_float((msb & ~31u) | 16u) : -FLT_MAX;
      const int64_t pair = (int64_t)bx * args.By + by;
      if (live && t < args.Tx) args.argX[pair * args.Tx + t] = (int16_t)mcol;
      float v = (live && t < lx) ? m : 0.f;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane == 0 && h * 4 + q < args.npX) args.partX[pair * args.npX + h * 4 + q] = v;
      reset(st);
    }
  } else {
    // ---- column finaliser: maximiser over all rows of the sample for every column of the tile ----
    int ct = 0;
    uint32_t keep5;
    asm("mov.u32 %0, 0xffffffe0;" : "=r"(keep5));
    for (int by = y0; by < y1; ++by) {
      const int ly = min(args.lenY[by], args.Ty);
      const int64_t pairY = (int64_t)by * args.Bx + bx;
      int16_t* ay = args.argY + pairY * args.Ty;
      float sum = 0.f;
      if (nh == 0) {
        // no valid row: the reference's amax over an all-masked axis is finfo.min for every valid column
        for (int s = lane; s < ly; s += 32) ay[s] = 0;
        for (int s = 0; s < ly; ++s) sum += -FLT_MAX;
      } else {
        for (int s0 = 0; s0 < ly; s0 += kTile, ++ct) {
          const int nb = ct & 1;
          mbar_wait(&c_full[nb], (uint32_t)(ct >> 1) & 1u);
          const uint32_t* buf = colbuf + (size_t)nb * kSlots * kTile;
          const int nvalid = min(ly - s0, kTile);
          // lane = columns 4 lane .. 4 lane + 3.  Slot keys -> (score bits | 15 - slot): one FMNMX3 tree over the 16
          // slots gives the winning slot (lowest slot = lowest rows on equal score bits); its key is fetched again
          // for the row inside the slot.  Slots of row tiles without a valid token hold kNegKey (set at start-up).
          float k[4][kSlots];
#pragma unroll
          for (int w = 0; w < kSlots; ++w) {
            const uint4 v = *reinterpret_cast<const uint4*>(buf + w * kTile + 4 * lane);
            const uint32_t vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              k[i][w] = __uint_as_float((vv[i] & keep5) | (uint32_t)(15 - w));
            }
          }
          int rows[4];
          float part = 0.f;
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float a = max3(max3(k[i][0], k[i][1], k[i][2]), max3(k[i][3], k[i][4], k[i][5]),
                                 max3(k[i][6], k[i][7], k[i][8]));
            const float b = max3(max3(k[i][9], k[i][10], k[i][11]), max3(k[i][12], k[i][13], k[i][14]), k[i][15]);
            const uint32_t best = __float_as_uint(fmaxf(a, b));
            const int slot = 15 - (int)(best & 15u);
            const uint32_t orig = buf[slot * kTile + 4 * lane + i];
            rows[i] = slot * 32 + 31 - (int)(orig & 31u);
            if (4 * lane + i < nvalid) part += __uint_as_float((best & ~31u) | 16u);
          }
          mbar_arrive(&c_empty[nb]);
          int16_t* dst = ay + s0 + 4 * lane;
          if (4 * lane + 3 < nvalid && (reinterpret_cast<uintptr_t>(dst) & 7) == 0) {
            uint2 pk;
            pk.x = (uint32_t)(rows[0] & 0xffff) | ((uint32_t)rows[1] << 16);
            pk.y = (uint32_t)(rows[2] & 0xffff) | ((uint32_t)rows[3] << 16);
            *reinterpret_cast<uint2*>(dst) = pk;
          } else {
#pragma unroll
            for (int i = 0; i < 4; ++i)
              if (4 * lane + i < nvalid) dst[i] = (int16_t)rows[i];
          }
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
          sum += part;
        }
      }
      if (lane == 0) args.partY[pairY] = sum;
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

// One sweep for both sides.  X must have Tx <= 512 (all of a sample's row tiles stay in shared memory).
//   outX[x * sx0 + y * sx1] = mean_t max_s S,  outY[y * sy0 + x * sy1] = mean_s max_t S
int tc_filip_both_fwd(const __nv_bfloat16* X, const __nv_bfloat16* Y, const int32_t* lenX, const int32_t* lenY, int64_t Bx,
                      int64_t By, int64_t Tx, int64_t Ty, const float* tau, float* partX, int npX, int16_t* argX,
                      float* partY, int16_t* argY, float* outX, int64_t sx0, int64_t sx1, float* outY, int64_t sy0,
                      int64_t sy1, cudaStream_t st) {
  PCLIP_REQUIRE(Tx <= f1::kMaxH * f1::kTile, PCLIP_ERR_UNSUPPORTED, "filip single sweep: Tx = %lld > 512", (long long)Tx);
  CUtensorMap mx, my;
  PCLIP_TRY(make_tmap_bf16_3d(&mx, X, (uint64_t)Bx, (uint64_t)Tx, f1::kDim, f1::kTile, 64));
  PCLIP_TRY(make_tmap_bf16_3d(&my, Y, (uint64_t)By, (uint64_t)Ty, f1::kDim, f1::kTile, 64));
  // y samples split over blockIdx.y so that the grid fills whole rounds of the SMs (one CTA per SM); a split costs the
  // CTA's fixed part (TMEM allocation, barriers, up to 128 KB of X tiles: about half a full-length sample pair).
  // cfg 3: 256 CTAs x 4 splits = 6.92 waves = 7 rounds of 64 pairs against 442.8 ideal.  The finest split within 1.5 %
  // of the best wins (ragged batches balance better).  PCLIP_FILIP_YSPLIT overrides.
  const int sms = std::max(device_sm_count(), 1);
  int ysplit = 1;
  {
    const int64_t cap = std::min<int64_t>(By, std::max<int64_t>(32, ceil_div(sms, Bx)));
    auto cost = [&](int64_t ys) { return (double)ceil_div(Bx * ys, (int64_t)sms) * ((double)ceil_div(By, ys) + 0.6); };
    double best = 1e300;
    for (int64_t ys = 1; ys <= cap; ++ys) best = std::min(best, cost(ys));
    for (int64_t ys = 1; ys <= cap; ++ys)
      if (cost(ys) <= best * 1.015) ysplit = (int)ys;
    static const int forced = [] {
      const char* e = getenv("PCLIP_FILIP_YSPLIT");
      return e ? atoi(e) : 0;
    }();
    if (forced >= 1) ysplit = (int)std::min<int64_t>(forced, By);
  }
  const int yps = (int)ceil_div(By, ysplit);
  ysplit = (int)ceil_div(By, yps);
  static const int dbg = [] {
    const char* e = getenv("PCLIP_FILIP_DBG");
    return e ? atoi(e) : 0;
  }();
  FilipBothArgs a{(int)Bx, (int)By, (int)Tx, (int)Ty, lenX, lenY, yps, partX, npX, argX, partY, argY};
  auto* kern = dbg == 1 ? tc_filip_both_kernel<1> : dbg == 2 ? tc_filip_both_kernel<2> : tc_filip_both_kernel<0>;
  PCLIP_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)f1::kSmem));
  dim3 grid((unsigned)Bx, (unsigned)ysplit);
  kern<<<grid, f1::kThreads, f1::kSmem, st>>>(mx, my, a);
  PCLIP_CHECK_LAUNCH();
  PCLIP_TRY(filip_finalize(partX, npX, lenX, (int)Bx, (int)By, (int)Tx, tau, outX, sx0, sx1, st));
  return filip_finalize(partY, 1, lenY, (int)By, (int)Bx, (int)Ty, tau, outY, sy0, sy1, st);
}

}  // namespace pclip
