// CLIP InfoNCE forward + backward for SMALL batches (B <= 512, D == 128) in ONE cooperative launch.
//   ref: modules/training_utils.py:247-265 (_contrastive_loss_gc on 16 + 16 GradCache leaves of [16, 128]) and :74-89.
//
// configs[1] -- the reference's default training configuration (main.py:34,49: 16 x 16 accumulation) -- is 50 MFLOP:
// the large-batch pipeline (pack, two-sided sum-exp, finalize, gradient = 4 launches, then autograd scaling the
// gradients and handing 32 views to 32 AccumulateGrad nodes) spends 380 us on 40 us of GPU work.  This kernel reads the
// leaves through their pointers (no torch.cat, no pack), keeps the strip of exp(S) it owns in shared memory across a
// grid-wide barrier, and writes every leaf's gradient -- already multiplied by 1/(2B) and by the device-side
// grad_output (GradScaler's scale) -- straight into the caller's per-leaf buffers.
//
//   CTA = 8 rows of one modality ("strip"; 2 B / 8 CTAs, all resident: cooperative launch), 128 threads.
//   phase A  s[8, B] = a_strip . b^T in fp32 (b staged 64 rows at a time), e = exp(s) kept in shared memory,
//            row sums of e and the diagonal logits to global                      -> grid barrier
//   phase B  loss (CTA 0); g = e (1/R_i + 1/C_j) - 2 [j == i], dA_strip = g . b (second pass over the staged b rows),
//            scaled and stored through the leaf pointers
// Both modalities run the same code with the roles of x and y swapped: R of one side is C of the other, so every
// gradient row has exactly one owner (no atomics, deterministic).  Arithmetic is fp32 FFMA: at this size the tensor
// cores would save ~2 us of 10 and cost the bf16 rounding.
#include <cooperative_groups.h>
#include <cuda_bf16.h>
#include <math.h>

#include "common.cuh"

namespace cg = cooperative_groups;

namespace pclip {

namespace cs {
constexpr int kDim = 128;
constexpr int kStrip = 8;        // rows per CTA
constexpr int kThreads = 128;
constexpr int kTile = 64;        // rows of the other modality staged per step
constexpr int kLd = kDim + 4;    // padded row length in shared memory (132 = 4 mod 32: conflict-free 16-byte reads)
constexpr int kMaxChunks = 32;
constexpr int kMaxB = 512;

struct Side {
  const void* in[kMaxChunks];    // leaf data
  float* out[kMaxChunks];        // leaf gradient (fp32, same shape), may be null when no gradient is wanted
  int32_t row_begin[kMaxChunks + 1];
  int64_t in_rs[kMaxChunks], in_cs[kMaxChunks];     // element strides of the leaf
  int64_t out_rs[kMaxChunks];                        // row stride of the gradient buffer (contiguous columns)
  int n;
};

struct Args {
  Side side[2];          // [0] = x, [1] = y
  int B;
  int in_bf16;
  const float* gscale;   // device scalar multiplied into the gradients (may be null)
  float* loss;           // [1]
  float* rowsum;         // [2][B] workspace: sum_j exp(s_ij) per row of each side
  float* diag;           // [B]    workspace: s_ii
  int want_grad;
  int vec;               // every leaf is fp32, row-major (col stride 1), 16-byte aligned rows: float4 staging
};

__device__ __forceinline__ float load_elem(const Side& s, int chunk, int row_in_chunk, int d, int in_bf16) {
  const int64_t off = (int64_t)row_in_chunk * s.in_rs[chunk] + (int64_t)d * s.in_cs[chunk];
  return in_bf16 ? __bfloat162float(static_cast<const __nv_bfloat16*>(s.in[chunk])[off])
                 : static_cast<const float*>(s.in[chunk])[off];
}

__device__ __forceinline__ int find_chunk(const Side& s, int row) {
  int k = 0;
  while (k + 1 < s.n && row >= s.row_begin[k + 1]) ++k;
  return k;
}

// rows [r0, r0 + nrows) of `s` -> dst[nrows][kLd] (zero-filled past B); one warp per row, lane = 4 dims
__device__ __forceinline__ void stage_rows(const Side& s, int r0, int nrows, int B, int in_bf16, float* dst) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int r = warp; r < nrows; r += kThreads / 32) {
    const int row = r0 + r;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (row < B) {
      const int k = find_chunk(s, row);
      const int rr = row - s.row_begin[k];
      v.x = load_elem(s, k, rr, lane * 4, in_bf16);
      v.y = load_elem(s, k, rr, lane * 4 + 1, in_bf16);
      v.z = load_elem(s, k, rr, lane * 4 + 2, in_bf16);
      v.w = load_elem(s, k, rr, lane * 4 + 3, in_bf16);
    }
    *reinterpret_cast<float4*>(dst + r * kLd + lane * 4) = v;
  }
}
// Vector path: the 64 rows of a tile are 16 independent float4 loads per thread, issued together (and, in the loops
// below, one tile ahead of the arithmetic): a row-by-row staging loop paid one global-memory latency per row and made
// the kernel 110 us long at B = 256 (profiles/r02_clip_small_fused_kernel_ncu_full.txt: 5 % issue utilisation).
__device__ __forceinline__ void load_tile(const float* const* rowp, int j0, int B, float4 (&pre)[kTile / 4]) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
  for (int it = 0; it < kTile / 4; ++it) {
    const int row = j0 + it * 4 + warp;
    pre[it] = row < B ? __ldg(reinterpret_cast<const float4*>(rowp[row]) + lane) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
}

__device__ __forceinline__ void store_tile(const float4 (&pre)[kTile / 4], float* b_s) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
  for (int it = 0; it < kTile / 4; ++it) *reinterpret_cast<float4*>(b_s + (it * 4 + warp) * kLd + lane * 4) = pre[it];
}
}  // namespace cs

__global__ void __launch_bounds__(cs::kThreads) clip_small_fused_kernel(const __grid_constant__ cs::Args args) {
  using namespace cs;
  extern __shared__ __align__(16) float smem[];
  const int B = args.B;
  const int Bpad = (B + kTile - 1) / kTile * kTile + 4;   // + 4: the two strip rows a warp reads sit in different banks
  float* a_s = smem;                        // [kStrip][kLd]   this CTA's rows
  float* b_s = a_s + kStrip * kLd;          // [kTile][kLd]    staged rows of the other modality
  float* e_s = b_s + kTile * kLd;           // [kStrip][Bpad]  exp(s), later g
  float* red = e_s + kStrip * Bpad;         // [kStrip][16]    row-sum partials
  const float** rowp = reinterpret_cast<const float**>(red + kStrip * 16);   // [B] row starts of the other modality
  const int strips = (B + kStrip - 1) / kStrip;
  const int sd = blockIdx.x / strips;       // 0: rows of x against y, 1: rows of y against x
  const int r0 = (blockIdx.x - sd * strips) * kStrip;
  const cs::Side& A = args.side[sd];
  const cs::Side& Bo = args.side[1 - sd];
  const int tid = threadIdx.x;
  const int r = tid >> 4, cgp = tid & 15;   // phase A: row r of the strip, columns cgp + 16 k of the staged tile

  const bool vec = args.vec != 0;
  if (vec) {
    for (int row = tid; row < B; row += kThreads) {
      const int k = find_chunk(Bo, row);
      rowp[row] = static_cast<const float*>(Bo.in[k]) + (int64_t)(row - Bo.row_begin[k]) * Bo.in_rs[k];
    }
  }
  stage_rows(A, r0, kStrip, B, args.in_bf16, a_s);
  __syncthreads();
  float4 pre[kTile / 4];
  if (vec) load_tile(rowp, 0, B, pre);
  float rsum = 0.f;
  for (int j0 = 0; j0 < B; j0 += kTile) {
    __syncthreads();                        // previous tile consumed
    if (vec) store_tile(pre, b_s); else stage_rows(Bo, j0, kTile, B, args.in_bf16, b_s);
    __syncthreads();
    if (vec && j0 + kTile < B) load_tile(rowp, j0 + kTile, B, pre);      // in flight under this tile's arithmetic
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 8
    for (int d = 0; d < kDim; d += 4) {
      const float4 av = *reinterpret_cast<const float4*>(a_s + r * kLd + d);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float4 bv = *reinterpret_cast<const float4*>(b_s + (cgp + 16 * k) * kLd + d);
        acc[k] = fmaf(av.x, bv.x, acc[k]);
        acc[k] = fmaf(av.y, bv.y, acc[k]);
        acc[k] = fmaf(av.z, bv.z, acc[k]);
        acc[k] = fmaf(av.w, bv.w, acc[k]);
      }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int j = j0 + cgp + 16 * k;
      const bool ok = j < B && r0 + r < B;
      const float e = ok ? expf(acc[k]) : 0.f;        // the reference's plain exp (no max subtraction), :86
      e_s[r * Bpad + j] = e;
      rsum += e;
      if (ok && sd == 0 && j == r0 + r) args.diag[j] = acc[k];
    }
  }
  red[r * 16 + cgp] = rsum;
  __syncthreads();
  if (tid < kStrip && r0 + tid < B) {
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 16; ++k) s += red[tid * 16 + k];   // fixed order: deterministic
    args.rowsum[sd * B + r0 + tid] = s;
  }
  __threadfence();
  cg::this_grid().sync();

  // ---- phase B ----
  const float* Ra = args.rowsum + sd * B;          // normalisers of this side's rows ...
  const float* Cb = args.rowsum + (1 - sd) * B;    // ... and of the other side's rows (= the columns of this strip)
  if (blockIdx.x == 0) {
    // loss = 1/(2B) sum_i (log R_i + log C_i - 2 s_ii), one warp, fixed order
    if (tid < 32) {
      float l = 0.f;
      for (int i = tid; i < B; i += 32) l += logf(args.rowsum[i]) + logf(args.rowsum[B + i]) - 2.f * args.diag[i];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) l += __shfl_xor_sync(0xffffffffu, l, o);
      if (tid == 0) args.loss[0] = l * (0.5f / (float)B);
    }
  }
  if (!args.want_grad) return;
  // g = e (1/R_i + 1/C_j) - 2 [j == i]     (ref: dS of :85-89, without the 1/(2B) factor)
  for (int idx = tid; idx < kStrip * Bpad; idx += kThreads) {
    const int rr = idx / Bpad, j = idx - rr * Bpad;
    const int i = r0 + rr;
    float g = 0.f;
    if (i < B && j < B) g = e_s[idx] * (1.f / Ra[i] + 1.f / Cb[j]) - (j == i ? 2.f : 0.f);
    e_s[idx] = g;
  }
  // dA[r, :] = sum_j g[r, j] b_j: thread = (row r, dims dq*4 .. +3 and 64 + dq*4 .. +3)
  const int dq = tid & 15;
  float4 d0 = make_float4(0.f, 0.f, 0.f, 0.f), d1 = d0;
  if (vec) load_tile(rowp, 0, B, pre);
  for (int j0 = 0; j0 < B; j0 += kTile) {
    __syncthreads();
    if (vec) store_tile(pre, b_s); else stage_rows(Bo, j0, kTile, B, args.in_bf16, b_s);
    __syncthreads();
    if (vec && j0 + kTile < B) load_tile(rowp, j0 + kTile, B, pre);
#pragma unroll 8
    for (int j = 0; j < kTile; ++j) {
      const float g = e_s[r * Bpad + j0 + j];
      const float4 b0 = *reinterpret_cast<const float4*>(b_s + j * kLd + dq * 4);
      const float4 b1 = *reinterpret_cast<const float4*>(b_s + j * kLd + 64 + dq * 4);
      d0.x = fmaf(g, b0.x, d0.x), d0.y = fmaf(g, b0.y, d0.y), d0.z = fmaf(g, b0.z, d0.z), d0.w = fmaf(g, b0.w, d0.w);
      d1.x = fmaf(g, b1.x, d1.x), d1.y = fmaf(g, b1.y, d1.y), d1.z = fmaf(g, b1.z, d1.z), d1.w = fmaf(g, b1.w, d1.w);
    }
  }
  const int row = r0 + r;
  if (row < B) {
    const float sc = (0.5f / (float)B) * (args.gscale ? args.gscale[0] : 1.f);
    const int k = find_chunk(A, row);
    if (A.out[k]) {
      float* dst = A.out[k] + (int64_t)(row - A.row_begin[k]) * A.out_rs[k];
      *reinterpret_cast<float4*>(dst + dq * 4) = make_float4(d0.x * sc, d0.y * sc, d0.z * sc, d0.w * sc);
      *reinterpret_cast<float4*>(dst + 64 + dq * 4) = make_float4(d1.x * sc, d1.y * sc, d1.z * sc, d1.w * sc);
    }
  }
}

}  // namespace pclip

using namespace pclip;

static int fill_side(cs::Side& s, const pclip_chunk_t* in, const pclip_chunk_t* out, int n, int64_t B, int in_dtype) {
  PCLIP_REQUIRE(in && n > 0 && n <= cs::kMaxChunks, PCLIP_ERR_UNSUPPORTED, "clip_small: 1 .. %d chunks per modality",
                cs::kMaxChunks);
  s.n = n;
  int64_t r = 0;
  for (int i = 0; i < n; ++i) {
    PCLIP_REQUIRE(in[i].ptr && in[i].rows >= 0, PCLIP_ERR_INVALID, "clip_small: bad chunk %d", i);
    s.in[i] = in[i].ptr;
    s.in_rs[i] = in[i].row_stride;
    s.in_cs[i] = in[i].col_stride;
    s.row_begin[i] = (int32_t)r;
    s.out[i] = nullptr;
    s.out_rs[i] = 0;
    if (out) {
      PCLIP_REQUIRE(out[i].ptr && out[i].rows == in[i].rows && out[i].col_stride == 1 &&
                        (reinterpret_cast<uintptr_t>(out[i].ptr) & 15) == 0 && (out[i].row_stride & 3) == 0,
                    PCLIP_ERR_INVALID, "clip_small: gradient chunk %d must match its leaf, be 16-byte aligned and row-major", i);
      s.out[i] = static_cast<float*>(const_cast<void*>(out[i].ptr));
      s.out_rs[i] = out[i].row_stride;
    }
    r += in[i].rows;
  }
  s.row_begin[n] = (int32_t)r;
  PCLIP_REQUIRE(r == B, PCLIP_ERR_INVALID, "clip_small: chunk rows sum to %lld, expected %lld", (long long)r, (long long)B);
  (void)in_dtype;
  return PCLIP_OK;
}

extern "C" {

int32_t pclip_clip_small_max_batch(void) { return cs::kMaxB; }

size_t pclip_clip_small_workspace_bytes(int64_t B) { return align_up((size_t)3 * B * sizeof(float), 1024); }

int pclip_clip_small_fused(const pclip_chunk_t* x_chunks, int32_t nx, const pclip_chunk_t* y_chunks, int32_t ny,
                           int32_t in_dtype, int64_t B, int32_t D, const float* gscale, float* loss,
                           const pclip_chunk_t* dx_chunks, const pclip_chunk_t* dy_chunks, void* workspace,
                           size_t workspace_bytes, pclip_stream_t stream) {
  PCLIP_TRY(require_device());
  PCLIP_REQUIRE(loss && B > 0, PCLIP_ERR_INVALID, "clip_small: null argument");
  PCLIP_REQUIRE(D == cs::kDim && B <= cs::kMaxB, PCLIP_ERR_UNSUPPORTED,
                "clip_small: needs D == %d and B <= %d (got D = %d, B = %lld); use pclip_clip_loss", cs::kDim, cs::kMaxB, D,
                (long long)B);
  PCLIP_REQUIRE((dx_chunks == nullptr) == (dy_chunks == nullptr), PCLIP_ERR_INVALID, "clip_small: dx and dy go together");
  PCLIP_REQUIRE(in_dtype == PCLIP_F32 || in_dtype == PCLIP_BF16, PCLIP_ERR_UNSUPPORTED, "clip_small: unknown dtype");
  PCLIP_REQUIRE(workspace && (reinterpret_cast<uintptr_t>(workspace) & 1023) == 0 &&
                    workspace_bytes >= pclip_clip_small_workspace_bytes(B),
                PCLIP_ERR_WORKSPACE, "clip_small: workspace missing, misaligned or too small");
  cs::Args a;
  PCLIP_TRY(fill_side(a.side[0], x_chunks, dx_chunks, nx, B, in_dtype));
  PCLIP_TRY(fill_side(a.side[1], y_chunks, dy_chunks, ny, B, in_dtype));
  a.B = (int)B;
  a.in_bf16 = in_dtype == PCLIP_BF16;
  a.gscale = gscale;
  a.loss = loss;
  a.rowsum = static_cast<float*>(workspace);
  a.diag = a.rowsum + 2 * B;
  a.want_grad = dx_chunks != nullptr;
  a.vec = in_dtype == PCLIP_F32;
  for (int sd = 0; sd < 2 && a.vec; ++sd)
    for (int i = 0; i < a.side[sd].n; ++i)
      if (a.side[sd].in_cs[i] != 1 || (a.side[sd].in_rs[i] & 3) != 0 || (reinterpret_cast<uintptr_t>(a.side[sd].in[i]) & 15) != 0)
        a.vec = 0;
  const int strips = (int)ceil_div(B, cs::kStrip);
  const int Bpad = (int)ceil_div(B, cs::kTile) * cs::kTile + 4;
  const size_t smem = sizeof(float) * ((size_t)cs::kStrip * cs::kLd + (size_t)cs::kTile * cs::kLd +
                                       (size_t)cs::kStrip * Bpad + cs::kStrip * 16) + sizeof(void*) * (size_t)B + 16;
  PCLIP_CHECK_CUDA(cudaFuncSetAttribute(clip_small_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int per_sm = 0;
  PCLIP_CHECK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, clip_small_fused_kernel, cs::kThreads, smem));
  PCLIP_REQUIRE(2 * strips <= per_sm * device_sm_count(), PCLIP_ERR_UNSUPPORTED,
                "clip_small: %d CTAs do not fit one cooperative wave on this device", 2 * strips);
  void* params[] = {&a};
  PCLIP_CHECK_CUDA(cudaLaunchCooperativeKernel(reinterpret_cast<void*>(clip_small_fused_kernel), dim3(2 * strips),
                                               dim3(cs::kThreads), params, smem, static_cast<cudaStream_t>(stream)));
  count_launch();
  return PCLIP_OK;
}

}  // extern "C"
