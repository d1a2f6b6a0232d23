// FILIP backward for the bf16 path as a true gather / scatter over the stored argmax indices.
//   ref: autograd of modules/models.py:332-351 (einsum -> masked_fill -> amax over s / over t -> masked means)
//
//   dX[x,t,:] = sum_y ( wX[x,y] * Y[y, argX[x,y,t], :]                       "gather":  row t's own maximiser
//                     + sum_{s : argY[y,x,s] == t} wY[y,x] * Y[y,s,:] )       "scatter": columns whose maximiser is t
//   wX = gX / (tau max(lenX,1e-6)),  wY = gY / (tau max(lenY,1e-6))
//
// The round-1 kernel ran this as a dense one-hot tensor-core contraction (2 T^2 D flop per sample pair for
// 4 D (lenX + lenY) algorithmic: 3.44 ms per side at cfg 3).  Here every non-zero costs one 256-byte row read and
// 128 FMAs:
//
//   filip_weights_kernel      (wX, wY) per sample pair, zero for empty samples
//   filip_invert_kernel       per (x, y): argY[y,x,:] -> entries (t << 16 | s) sorted by (t, s), and per group of RPW
//                             target rows the packed (first entry | count << 16).  Counting sort through a bit mask per
//                             target row, so the order -- and with it the fp32 summation order of the gradient -- is
//                             deterministic.  Persistent CTAs; the masks are cleared by their readers.
//   filip_gather_grad_kernel  warp = RPW consecutive token rows of sample x, lane = 4 embedding dims; the accumulators
//                             stay in registers over all y.  A gathered bf16 row is ONE coalesced 256-byte warp load
//                             from the L2-resident token matrix (33.5 MB at cfg 3).  Gather rows are static (row i of
//                             the warp); scatter entries arrive sorted by row and are dispatched through a switch, so
//                             the accumulators never leave the register file.  The kernel is bound by the latency of
//                             dependent loads (index -> row), so it runs many small warps (RPW = 4, 32 warps per SM)
//                             and fetches the index data of the next two sample pairs ahead of the current one.
#include <cuda_bf16.h>

#include "common.cuh"

namespace pclip {

namespace fb {
constexpr int kDim = 128;
constexpr int kInvThreads = 512;
constexpr int kInvRows = 512;                  // target rows per pass of the inversion
constexpr int kInvCols = 512;                  // source columns per pass (one bit each)
constexpr int kInvWords = kInvCols / 32;
}  // namespace fb

__host__ __device__ static inline int filip_groups(int64_t T, int rpw) { return (int)((T + rpw - 1) / rpw); }

// ---------------------------------------------------------------------------------------------------------
// (wX, wY) per pair
// ---------------------------------------------------------------------------------------------------------
__global__ void filip_weights_kernel(const int32_t* __restrict__ lenX, const int32_t* __restrict__ lenY, int Bx, int By,
                                     int Tx, int Ty, const float* __restrict__ tau, const float* __restrict__ gX,
                                     int64_t gx0, int64_t gx1, const float* __restrict__ gY, int64_t gy0, int64_t gy1,
                                     float2* __restrict__ w2) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= (int64_t)Bx * By) return;
  const int x = (int)(i / By), y = (int)(i % By);
  const int lx = min(lenX[x], Tx), ly = min(lenY[y], Ty);
  const float inv_tau = 1.0f / tau[0];
  float2 w = make_float2(0.f, 0.f);
  if (lx > 0 && ly > 0) {
    w.x = gX[x * gx0 + y * gx1] * inv_tau / fmaxf((float)lx, 1e-6f);
    w.y = gY[y * gy0 + x * gy1] * inv_tau / fmaxf((float)ly, 1e-6f);
  }
  w2[i] = w;
}

// ---------------------------------------------------------------------------------------------------------
// inversion: persistent CTAs of 512 threads over the Bx * By sample pairs
//   inv   [Bx][By][Ty]   (t << 16) | s for every s < lenY[y], sorted by (t, s)
//   grp2  [Bx][By][G]    per group g of `gs` target rows: first entry | (entry count << 16)   (G = ceil(Tx / gs))
// gs must divide 512.
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(fb::kInvThreads)
filip_invert_kernel(const int16_t* __restrict__ argY /* [By][Bx][Ty] */, const int32_t* __restrict__ lenX,
                    const int32_t* __restrict__ lenY, int Bx, int By, int Tx, int Ty, int gs,
                    uint32_t* __restrict__ inv, uint32_t* __restrict__ grp2) {
  using namespace fb;
  __shared__ uint32_t mask[kInvWords * kInvRows];   // 32 KB, [word][row] (conflict-free for the row's reader): bit (s - c0)
                                                    // of row (t - r0); cleared by the reader
  __shared__ int cnt[kInvRows + 1];                 // per row: entries so far (pass 1), start / write cursor (pass 2)
  __shared__ int warp_tot[kInvThreads / 32];
  __shared__ int range_base;
  __shared__ uint32_t summ[kInvRows];               // fast path: bit j = mask word j of the row is non-zero
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int G = filip_groups(Tx, gs);
  for (int i = tid; i < kInvRows * kInvWords; i += kInvThreads) mask[i] = 0u;
  summ[tid] = 0u;
  __syncthreads();
  const int gshift = 31 - __clz(gs);   // gs is a power of two
  for (int pair = blockIdx.x; pair < Bx * By; pair += gridDim.x) {   // Bx * By < 2^31 (checked by the host)
    const int x = pair / By, y = pair - x * By;
    const int lx = min(lenX[x], Tx), ly = min(lenY[y], Ty);
    const int16_t* ay = argY + ((int64_t)y * Bx + x) * Ty;
    uint32_t* out = inv + (int64_t)pair * Ty;
    uint32_t* go = grp2 + (int64_t)pair * G;
    if (ly <= 0 || lx <= 0) {   // CTA-uniform: no entries
      for (int g = tid; g < G; g += kInvThreads) go[g] = 0u;
      continue;
    }
    if (ly <= kInvCols && Tx <= kInvRows) {
      // ---- fast path (one column chunk, one row range): a per-row summary word says which mask words are set, so a
      //      row's thread touches only the ~1 words that hold its entries ----
      if (tid < ly) {
        const int r = (int)ay[tid];
        atomicOr(&mask[(tid >> 5) * kInvRows + r], 1u << (tid & 31));
        atomicOr(&summ[r], 1u << (tid >> 5));
      }
      __syncthreads();
      const uint32_t sm = summ[tid];
      int mine = 0;
      for (uint32_t m = sm; m; m &= m - 1) mine += __popc(mask[(__ffs(m) - 1) * kInvRows + tid]);
      int incl = mine;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
      }
      if (lane == 31) warp_tot[warp] = incl;
      __syncthreads();
      // exclusive prefix over the warps: lane k holds warp k's total
      int pre = lane < warp ? warp_tot[lane & 15] : 0;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) pre += __shfl_xor_sync(0xffffffffu, pre, o);
      const int start = pre + incl - mine;          // entries with t < tid
      cnt[tid] = start;
      if (tid == kInvThreads - 1) cnt[kInvRows] = start + mine;
      __syncthreads();
      if ((tid & (gs - 1)) == 0 && tid < Tx)
        go[tid >> gshift] = (uint32_t)start | ((uint32_t)(cnt[tid + gs] - start) << 16);
      if (sm) {
        summ[tid] = 0u;
        int pos = start;
        for (uint32_t m = sm; m; m &= m - 1) {
          const int j = __ffs(m) - 1;
          uint32_t w = mask[j * kInvRows + tid];
          mask[j * kInvRows + tid] = 0u;
          for (; w; w &= w - 1) out[pos++] = ((uint32_t)tid << 16) | (uint32_t)(j * 32 + __ffs(w) - 1);
        }
      }
      __syncthreads();   // a slow row owner may still be reading its mask words: close the pair before the next atomics
      continue;
    }
    if (tid == 0) range_base = 0;
    const int nchunk = (ly + kInvCols - 1) / kInvCols;
    for (int r0 = 0; r0 < Tx; r0 += kInvRows) {
      // ---- pass 1: entries per target row of this row range ----
      int mine = 0;
      uint32_t w[kInvWords];
      for (int c = 0; c < nchunk; ++c) {
        const int s = c * kInvCols + tid;
        if (s < ly) {
          const int r = (int)ay[s] - r0;
          if ((unsigned)r < (unsigned)kInvRows) atomicOr(&mask[(tid >> 5) * kInvRows + r], 1u << (tid & 31));
        }
        __syncthreads();
        int n = 0;
#pragma unroll
        for (int j = 0; j < kInvWords; ++j) {
          w[j] = mask[j * kInvRows + tid];
          n += __popc(w[j]);
        }
        mine += n;
        if (n) {   // the reader clears (multi-chunk: the mask is rebuilt in pass 2; single chunk: `w` is kept)
#pragma unroll
          for (int j = 0; j < kInvWords; ++j)
            if (w[j]) mask[j * kInvRows + tid] = 0u;
        }
        if (nchunk > 1) __syncthreads();
      }
      // ---- exclusive scan over the 512 rows of the range ----
      int incl = mine;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
      }
      if (lane == 31) warp_tot[warp] = incl;
      __syncthreads();
      int pre = range_base;
      for (int k = 0; k < warp; ++k) pre += warp_tot[k];
      const int start = pre + incl - mine;          // entries with t < r0 + tid
      const int t = r0 + tid;
      cnt[tid] = start;
      if (tid == kInvThreads - 1) cnt[kInvRows] = start + mine;
      __syncthreads();
      if (tid == 0) range_base = cnt[kInvRows];
      if ((tid & (gs - 1)) == 0 && t < Tx) go[t >> gshift] = (uint32_t)start | ((uint32_t)(cnt[tid + gs] - start) << 16);
      // ---- pass 2: place the entries ----
      if (nchunk == 1) {
        if (mine) {
          int pos = start;
#pragma unroll
          for (int j = 0; j < kInvWords; ++j) {
            uint32_t m = w[j];
            while (m) {
              const int b = __ffs(m) - 1;
              m &= m - 1;
              out[pos++] = ((uint32_t)t << 16) | (uint32_t)(j * 32 + b);
            }
          }
        }
      } else {
        __syncthreads();   // cnt is turned into the write cursor below: group leaders have read it
        for (int c = 0; c < nchunk; ++c) {
          const int s = c * kInvCols + tid;
          if (s < ly) {
            const int r = (int)ay[s] - r0;
            if ((unsigned)r < (unsigned)kInvRows) atomicOr(&mask[(tid >> 5) * kInvRows + r], 1u << (tid & 31));
          }
          __syncthreads();
          int pos = cnt[tid];
#pragma unroll
          for (int j = 0; j < kInvWords; ++j) {
            uint32_t m = mask[j * kInvRows + tid];
            if (m) mask[j * kInvRows + tid] = 0u;
            while (m) {
              const int b = __ffs(m) - 1;
              m &= m - 1;
              out[pos++] = ((uint32_t)t << 16) | (uint32_t)(c * kInvCols + j * 32 + b);
            }
          }
          cnt[tid] = pos;
          __syncthreads();
        }
      }
      // the next range / pair starts with atomicOr into the (clean) mask and ends its first phase with a barrier
      // before anything read here (warp_tot, cnt, range_base) is rewritten
    }
  }
}

// ---------------------------------------------------------------------------------------------------------
// inversion, one WARP per sample pair (Tx, Ty <= 512, groups of 4 or 8 rows): no block barrier anywhere, so 48 warps per SM
// work on 48 pairs at once (the CTA-per-pair kernel above spends its time in __syncthreads: 0.56 ms per side at cfg 3).
// Stable counting sort of the pair's <= 512 maximisers by target row:
//   1. histogram of the target rows in a warp-private shared-memory array (the 16 coalesced index loads of the pair
//      are in flight together and stay in registers for step 3)
//   2. exclusive scan (lane = 16 consecutive rows) -> write cursors and the packed (first entry | count << 16) words
//   3. batches of 32 columns in column order: rank among the lanes of the batch with the same row (each ORs its bit into the row's shared-memory word), entry
//      to cursor + rank of a shared-memory image of the output; the first lane of each row advances the cursor
//   4. the image goes out with coalesced 16-byte stores
// The order -- (row, column) ascending -- is the one the CTA kernel produces, so the gradients stay bit-identical.
// ---------------------------------------------------------------------------------------------------------
namespace fb {
constexpr int kWInvWarps = 4;
constexpr int kWInvMax = 512;
}  // namespace fb

__global__ void __launch_bounds__(32 * fb::kWInvWarps, 8)
filip_invert_warp_kernel(const int16_t* __restrict__ argY /* [By][Bx][Ty] */, const int32_t* __restrict__ lenX,
                         const int32_t* __restrict__ lenY, int Bx, int By, int Tx, int Ty, int gs /* 4 or 8 */,
                         uint32_t* __restrict__ inv, uint32_t* __restrict__ grp2) {
  using namespace fb;
  __shared__ __align__(16) uint32_t cnt_s[kWInvWarps][kWInvMax];
  __shared__ __align__(16) uint32_t out_s[kWInvWarps][kWInvMax];
  __shared__ __align__(16) uint32_t bit_s[kWInvWarps][kWInvMax];   // per row: lanes of the current batch that target it
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t* cnt = cnt_s[warp];
  uint32_t* img = out_s[warp];
  uint32_t* bits = bit_s[warp];
#pragma unroll
  for (int i = 0; i < 4; ++i) *reinterpret_cast<uint4*>(bits + lane * 16 + i * 4) = make_uint4(0u, 0u, 0u, 0u);
  __syncwarp();
  const int G = filip_groups(Tx, gs);
  const uint32_t lt = (1u << lane) - 1u;
  for (int pair = blockIdx.x * kWInvWarps + warp; pair < Bx * By; pair += gridDim.x * kWInvWarps) {
    const int x = pair / By, y = pair - x * By;
    const int lx = min(lenX[x], Tx), ly = min(lenY[y], Ty);
    uint32_t* go = grp2 + (int64_t)pair * G;
    if (ly <= 0 || lx <= 0) {   // warp-uniform: no entries
      for (int g = lane; g < G; g += 32) go[g] = 0u;
      continue;
    }
    const int16_t* ay = argY + ((int64_t)y * Bx + x) * Ty;
    uint32_t* out = inv + (int64_t)pair * Ty;
    const int nb = (ly + 31) >> 5;
    uint32_t tv[kWInvMax / 32];
#pragma unroll
    for (int b = 0; b < kWInvMax / 32; ++b) {
      const int s = b * 32 + lane;
      tv[b] = (b < nb && s < ly) ? (uint32_t)(uint16_t)__ldg(ay + s) : 0xffffu;
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) *reinterpret_cast<uint4*>(cnt + lane * 16 + i * 4) = make_uint4(0u, 0u, 0u, 0u);
    __syncwarp();
#pragma unroll
    for (int b = 0; b < kWInvMax / 32; ++b)
      if (tv[b] != 0xffffu) atomicAdd(&cnt[tv[b]], 1u);
    __syncwarp();
    // exclusive scan: lane owns rows 16 lane .. 16 lane + 15
    uint32_t c[16];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const uint4 v = *reinterpret_cast<const uint4*>(cnt + lane * 16 + i * 4);
      c[4 * i] = v.x, c[4 * i + 1] = v.y, c[4 * i + 2] = v.z, c[4 * i + 3] = v.w;
    }
    uint32_t tot = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) tot += c[i];
    uint32_t incl = tot;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += v;
    }
    uint32_t run = incl - tot, first8 = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const uint32_t s0 = run, s1 = s0 + c[4 * i], s2 = s1 + c[4 * i + 1], s3 = s2 + c[4 * i + 2];
      run = s3 + c[4 * i + 3];
      *reinterpret_cast<uint4*>(cnt + lane * 16 + i * 4) = make_uint4(s0, s1, s2, s3);
      if (gs == 4) {
        if (lane * 4 + i < G) go[lane * 4 + i] = s0 | ((run - s0) << 16);
      } else if ((i & 1) == 0) {
        first8 = s0;
      } else if (lane * 2 + (i >> 1) < G) {
        go[lane * 2 + (i >> 1)] = first8 | ((run - first8) << 16);
      }
    }
    __syncwarp();
    // placement, batch by batch in column order (stable)
#pragma unroll
    for (int b = 0; b < kWInvMax / 32; ++b) {
      if (b < nb) {   // warp-uniform
        const uint32_t t = tv[b];
        const bool valid = t != 0xffffu;
        if (valid) atomicOr(&bits[t], 1u << lane);   // lanes of the batch with the same row claim a bit of its word
        __syncwarp();
        const uint32_t m = valid ? bits[t] : 0u;
        const uint32_t rank = __popc(m & lt);
        const uint32_t base = valid ? cnt[t] : 0u;
        if (valid) img[base + rank] = (t << 16) | (uint32_t)(b * 32 + lane);
        __syncwarp();
        if (valid && rank == 0) {   // the row's first lane advances the cursor and clears the word
          cnt[t] = base + __popc(m);
          bits[t] = 0u;
        }
        __syncwarp();
      }
    }
    if ((Ty & 3) == 0) {
      for (int i = lane * 4; i < ly; i += 128) *reinterpret_cast<uint4*>(out + i) = *reinterpret_cast<const uint4*>(img + i);
    } else {
      for (int i = lane; i < ly; i += 32) out[i] = img[i];
    }
    __syncwarp();
  }
}

// ---------------------------------------------------------------------------------------------------------
// gradient: grid ceil(Bx * ceil(Tx / RPW) / WARPS), 32 * WARPS threads, (Bx + 1) ints of dynamic shared memory
// ---------------------------------------------------------------------------------------------------------
struct FilipGatherArgs {
  int Bx, By, Tx, Ty;
  const int32_t* lenX;
  const float2* w2;              // [Bx][By] (wX, wY)
  const int16_t* argX;           // [Bx][By][Tx]
  const uint32_t* inv;           // [Bx][By][Ty]   entries of argY[y,x,:] sorted by target row
  const uint32_t* grp2;          // [Bx][By][G]    first entry | count << 16 per group of RPW rows
  const __nv_bfloat16* Yr;       // [By][Ty][128] bf16 token rows (pads zero)
  float* dX;                     // [Bx][Tx][128]
};

namespace fb {
// acc[0..3] += w * (4 bf16 of the row held in v): dims 4 lane .. 4 lane + 3.  Two packed fp32 FMAs (FFMA2, sm_100).
__device__ __forceinline__ void fma_row(float (&a)[4], const uint2 v, const float w) {
  uint64_t a01, a23, b01, b23, ww;
  asm("mov.b64 %0, {%1, %2};" : "=l"(a01) : "f"(a[0]), "f"(a[1]));
  asm("mov.b64 %0, {%1, %2};" : "=l"(a23) : "f"(a[2]), "f"(a[3]));
  asm("mov.b64 %0, {%1, %2};" : "=l"(b01) : "r"(v.x << 16), "r"(v.x & 0xffff0000u));
  asm("mov.b64 %0, {%1, %2};" : "=l"(b23) : "r"(v.y << 16), "r"(v.y & 0xffff0000u));
  asm("mov.b64 %0, {%1, %1};" : "=l"(ww) : "f"(w));
  asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(a01) : "l"(b01), "l"(ww));
  asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(a23) : "l"(b23), "l"(ww));
  asm("mov.b64 {%0, %1}, %2;" : "=f"(a[0]), "=f"(a[1]) : "l"(a01));
  asm("mov.b64 {%0, %1}, %2;" : "=f"(a[2]), "=f"(a[3]) : "l"(a23));
}

template <int RPW>
__device__ __forceinline__ void fma_dispatch(float (&acc)[RPW][4], const uint32_t row, const uint2 v, const float w) {
  static_assert(RPW == 4 || RPW == 8, "rows per warp");
  switch (row) {
    case 0: fma_row(acc[0], v, w); break;
    case 1: fma_row(acc[1], v, w); break;
    case 2: fma_row(acc[2], v, w); break;
    case 3: fma_row(acc[3], v, w); break;
    default:
      if constexpr (RPW == 8) {
        switch (row) {
          case 4: fma_row(acc[4], v, w); break;
          case 5: fma_row(acc[5], v, w); break;
          case 6: fma_row(acc[6], v, w); break;
          default: fma_row(acc[7], v, w); break;
        }
      }
      break;
  }
}
}  // namespace fb

// One sample pair for one warp with exactly N scatter entries (N is warp-uniform): no predication anywhere.  The row
// loads of the gather (vg, already issued by the caller) and of the scatter are all in flight before the first FMA.
template <int RPW, int N>
__device__ __forceinline__ void pair_body_impl(float (&acc)[RPW][4], const uint2 (&vg)[RPW], const uint32_t seg,
                                          const uint2* __restrict__ yb, const float wx, const float wy) {
  using namespace fb;
  uint32_t e[N > 0 ? N : 1];
  uint2 v[N > 0 ? N : 1];
#pragma unroll
  for (int j = 0; j < N; ++j) {
    e[j] = __shfl_sync(0xffffffffu, seg, j);
    v[j] = __ldg(yb + (e[j] & 0xffffu) * (uint32_t)(kDim / 4));
  }
#pragma unroll
  for (int i = 0; i < RPW; ++i) fma_row(acc[i], vg[i], wx);
#pragma unroll
  for (int j = 0; j < N; ++j) fma_dispatch<RPW>(acc, (e[j] >> 16) & (RPW - 1), v[j], wy);
}

template <int RPW, int WARPS>
__global__ void __launch_bounds__(32 * WARPS, 1) filip_gather_grad_kernel(const FilipGatherArgs args) {
  using namespace fb;
  constexpr int kMaxN = 8;   // entries handled by the specialised bodies; more (rare) go through the tail loop
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;
  const int G = (args.Tx + RPW - 1) / RPW;

  // Work items = groups of RPW token rows.  Ragged batches: the groups that hold a valid token are numbered first
  // (sample by sample), the all-pad groups after them, and warp w of the grid takes item w -- so every warp of the
  // early CTAs has real work (a (row block, sample) grid left the warps past a short sample's length idle: 1.47 ms
  // against 1.97 ms for 0.44x the rows) and the late CTAs only write zeros.  pref[x] = valid groups of samples < x.
  extern __shared__ int pref[];
  {
    __shared__ int wtot[WARPS];
    int carry = 0;
    for (int base = 0; base < args.Bx; base += 32 * WARPS) {
      const int x = base + (int)threadIdx.x;
      const int v = x < args.Bx ? (min(args.lenX[x], args.Tx) + RPW - 1) / RPW : 0;
      int inc = v;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int nb = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += nb;
      }
      if (lane == 31) wtot[threadIdx.x >> 5] = inc;
      __syncthreads();
      int woff = 0, tot = 0;
      for (int w = 0; w < WARPS; ++w) {
        if (w < (int)(threadIdx.x >> 5)) woff += wtot[w];
        tot += wtot[w];
      }
      if (x < args.Bx) pref[x] = carry + woff + inc - v;
      carry += tot;
      __syncthreads();
    }
    if (threadIdx.x == 0) pref[args.Bx] = carry;
    __syncthreads();
  }
  const int total = pref[args.Bx];
  const int item = blockIdx.x * WARPS + warp;
  if (item >= args.Bx * G) return;
  int bx, wrow0;
  {
    // valid item: largest x with pref[x] <= item; pad item k = item - total: the same search on x * G - pref[x]
    const bool pad = item >= total;
    const int key = pad ? item - total : item;
    int lo = 0, hi = args.Bx;                       // invariant: f(lo) <= key < f(hi), f(Bx) = count of that kind
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      const int f = pad ? mid * G - pref[mid] : pref[mid];
      if (f <= key) lo = mid; else hi = mid;
    }
    bx = lo;
    const int valid_groups = pref[bx + 1] - pref[bx];
    wrow0 = (pad ? valid_groups + (key - (bx * G - pref[bx])) : key - pref[bx]) * RPW;
  }
  const int lx = min(args.lenX[bx], args.Tx);

  float acc[RPW][4];
#pragma unroll
  for (int i = 0; i < RPW; ++i) acc[i][0] = acc[i][1] = acc[i][2] = acc[i][3] = 0.f;

  if (wrow0 < lx) {   // warp-uniform: this warp owns at least one valid token
    const int trow = wrow0 + (lane & (RPW - 1));
    const bool row_ok = trow < lx;        // rows past lx of a partly valid group gather row 0 and are zeroed at the end
    const int64_t pair0 = (int64_t)bx * args.By;
    // running pointers, advanced by one sample pair per iteration
    const float2* pw = args.w2 + pair0;
    const uint32_t* pg = args.grp2 + pair0 * G + wrow0 / RPW;
    const int16_t* pa = args.argX + pair0 * args.Tx + trow;
    const uint32_t* pe = args.inv + pair0 * args.Ty + lane;
    const uint2* yb = reinterpret_cast<const uint2*>(args.Yr) + lane;
    const int ystep = args.Ty * (kDim / 4);
    // index data is fetched two pairs ahead (weights, entry range, maximisers) and one pair ahead (the entries)
    float2 w1 = __ldg(pw), w2 = __ldg(pw + (args.By > 1));
    uint32_t g1 = __ldg(pg), g2 = __ldg(pg + (args.By > 1 ? G : 0));
    int a1 = row_ok ? (int)__ldg(pa) : 0, a2 = row_ok ? (int)__ldg(pa + (args.By > 1 ? args.Tx : 0)) : 0;
    uint32_t s1 = lane < (int)(g1 >> 16) ? __ldg(pe + (g1 & 0xffffu)) : 0u;
    for (int by = 0; by < args.By; ++by) {
      const float2 w = w1;
      const uint32_t g = g1, seg = s1;
      const int a = a1;
      w1 = w2, g1 = g2, a1 = a2;
      // pair by + 1: entries; pair by + 2: level-1 data (clamped to the last pair: harmless repeats)
      const int d1 = by + 1 < args.By, d2 = by + 2 < args.By ? 2 : d1;
      s1 = lane < (int)(g1 >> 16) ? __ldg(pe + (size_t)d1 * args.Ty + (g1 & 0xffffu)) : 0u;
      w2 = __ldg(pw + d2);
      g2 = __ldg(pg + (size_t)d2 * G);
      a2 = row_ok ? (int)__ldg(pa + (size_t)d2 * args.Tx) : 0;
      const int n = (int)(g >> 16);
      uint2 vg[RPW];
#pragma unroll
      for (int i = 0; i < RPW; ++i) vg[i] = __ldg(yb + (uint32_t)__shfl_sync(0xffffffffu, a, i) * (uint32_t)(kDim / 4));
      switch (min(n, kMaxN)) {
        case 0: pair_body_impl<RPW, 0>(acc, vg, seg, yb, w.x, w.y); break;
        case 1: pair_body_impl<RPW, 1>(acc, vg, seg, yb, w.x, w.y); break;
        case 2: pair_body_impl<RPW, 2>(acc, vg, seg, yb, w.x, w.y); break;
        case 3: pair_body_impl<RPW, 3>(acc, vg, seg, yb, w.x, w.y); break;
        case 4: pair_body_impl<RPW, 4>(acc, vg, seg, yb, w.x, w.y); break;
        case 5: pair_body_impl<RPW, 5>(acc, vg, seg, yb, w.x, w.y); break;
        case 6: pair_body_impl<RPW, 6>(acc, vg, seg, yb, w.x, w.y); break;
        case 7: pair_body_impl<RPW, 7>(acc, vg, seg, yb, w.x, w.y); break;
        default: pair_body_impl<RPW, 8>(acc, vg, seg, yb, w.x, w.y); break;
      }
      if (n > kMaxN) {   // tokens of x that are the maximiser of many columns of y (short x against long y)
        const uint32_t* ent = pe - lane + (g & 0xffffu);
        uint32_t sg = seg;
        // four entries per step: their row loads are in flight together; a slot past n reads row 0 with weight 0
        for (int k = kMaxN; k < n; k += 4) {
          if ((k & 31) == 0) sg = k + lane < n ? __ldg(ent + k + lane) : 0u;     // kMaxN and the step divide 32
          uint32_t e[4];
          uint2 v[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            e[j] = __shfl_sync(0xffffffffu, sg, (k + j) & 31);
            v[j] = __ldg(yb + (e[j] & 0xffffu) * (uint32_t)(kDim / 4));
          }
#pragma unroll
          for (int j = 0; j < 4; ++j) fma_dispatch<RPW>(acc, (e[j] >> 16) & (RPW - 1), v[j], k + j < n ? w.y : 0.f);
        }
      }
      pw += 1, pg += G, pa += args.Tx, pe += args.Ty, yb += ystep;
    }
  }
  // rows [lx, Tx) receive exactly zero (no index ever points at a pad)
#pragma unroll
  for (int i = 0; i < RPW; ++i) {
    const int t = wrow0 + i;
    if (t < args.Tx)
      *reinterpret_cast<float4*>(args.dX + ((int64_t)bx * args.Tx + t) * kDim + lane * 4) =
          t < lx ? make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
}

// ---------------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------------
static int gather_rpw() {
  static int rpw = 0;
  if (!rpw) {
    const char* e = getenv("PCLIP_FILIP_RPW");   // A/B switch: 4 (default) or 8 token rows per warp
    rpw = (e && e[0] == '8') ? 8 : 4;
  }
  return rpw;
}

size_t filip_gather_inv_elems(int64_t Ba, int64_t Bb, int64_t TA, int64_t TB) {
  return (size_t)Ba * Bb * std::max(TA, TB);
}
size_t filip_gather_grp_elems(int64_t Ba, int64_t Bb, int64_t TA, int64_t TB) {
  return (size_t)Ba * Bb * filip_groups(std::max(TA, TB), 4);
}

// one side: dX for the samples of X against all samples of Y
int filip_gather_side(const __nv_bfloat16* Yr, const int32_t* lenX, const int32_t* lenY, int64_t Bx, int64_t By,
                      int64_t Tx, int64_t Ty, const float* tau, const float* gX, int64_t gx0, int64_t gx1,
                      const float* gY, int64_t gy0, int64_t gy1, const int16_t* argX, const int16_t* argY, float2* w2,
                      uint32_t* inv, uint32_t* grp2, float* dX, cudaStream_t st) {
  PCLIP_REQUIRE(Tx <= 32767 && Ty <= 32767, PCLIP_ERR_UNSUPPORTED, "filip_bwd: at most 32767 tokens per sequence");
  PCLIP_REQUIRE(Bx * By < (int64_t)1 << 31, PCLIP_ERR_UNSUPPORTED, "filip_bwd: more than 2^31 sample pairs");
  const int rpw = gather_rpw();
  filip_weights_kernel<<<(unsigned)ceil_div(Bx * By, 256), 256, 0, st>>>(lenX, lenY, (int)Bx, (int)By, (int)Tx, (int)Ty,
                                                                         tau, gX, gx0, gx1, gY, gy0, gy1, w2);
  PCLIP_CHECK_LAUNCH();
  static const bool cta_invert = [] {
    const char* e = getenv("PCLIP_FILIP_INVERT");   // A/B switch: "cta" forces the CTA-per-pair inversion
    return e && e[0] == 'c';
  }();
  if (!cta_invert && Tx <= fb::kWInvMax && Ty <= fb::kWInvMax) {
    const int64_t blocks = std::min<int64_t>(ceil_div(Bx * By, (int64_t)fb::kWInvWarps), (int64_t)device_sm_count() * 8);
    filip_invert_warp_kernel<<<(unsigned)blocks, 32 * fb::kWInvWarps, 0, st>>>(argY, lenX, lenY, (int)Bx, (int)By,
                                                                               (int)Tx, (int)Ty, rpw, inv, grp2);
  } else {
    const int64_t inv_ctas = std::min<int64_t>(Bx * By, (int64_t)device_sm_count() * 4);
    filip_invert_kernel<<<(unsigned)inv_ctas, fb::kInvThreads, 0, st>>>(argY, lenX, lenY, (int)Bx, (int)By, (int)Tx,
                                                                        (int)Ty, rpw, inv, grp2);
  }
  PCLIP_CHECK_LAUNCH();
  FilipGatherArgs a{(int)Bx, (int)By, (int)Tx, (int)Ty, lenX, w2, argX, inv, grp2, Yr, dX};
  PCLIP_REQUIRE(Bx <= 8192, PCLIP_ERR_UNSUPPORTED, "filip_bwd: at most 8192 samples per side (got %lld)", (long long)Bx);
  const size_t pref_bytes = (size_t)(Bx + 1) * sizeof(int);
  if (rpw == 4) {
    const unsigned gg = (unsigned)ceil_div(Bx * filip_groups(Tx, 4), (int64_t)32);
    filip_gather_grad_kernel<4, 32><<<gg, 1024, pref_bytes, st>>>(a);
  } else {
    const unsigned gg = (unsigned)ceil_div(Bx * filip_groups(Tx, 8), (int64_t)16);
    filip_gather_grad_kernel<8, 16><<<gg, 512, pref_bytes, st>>>(a);
  }
  PCLIP_CHECK_LAUNCH();
  return PCLIP_OK;
}

}  // namespace pclip
