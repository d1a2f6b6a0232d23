// HBM-bound passes around the contractions: chunk packing, L2-normalise (+ temperature), loss finalisation,
// scalar grad_output scaling and the row-wise InfoNCE tail of FILIP.  One warp per row, float4 accesses.
#include <cuda_bf16.h>
#include <float.h>

#include <mutex>

#include "common.cuh"

namespace pclip {

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// ---------------------------------------------------------------------------------------------
// pack: list of strided row chunks -> contiguous [rows, D] (fp32 or bf16), optionally also the transposed
// copy outT [D, ldT].  One launch packs up to two operands (blockIdx.y), so the CLIP composite packs x and y
// (GradCache's two lists of leaves, ref: modules/training_utils.py:247-249) with a single kernel.
// ---------------------------------------------------------------------------------------------
constexpr int kMaxChunks = 32;
struct PackOperand {
  const void* ptr[kMaxChunks];
  int64_t row_begin[kMaxChunks + 1];
  int64_t row_stride[kMaxChunks];
  int64_t col_stride[kMaxChunks];
  void* out;
  void* outT;
  int64_t ldT;
  int n;
};
struct PackArgs {
  PackOperand op[2];
};

template <typename T>
__device__ __forceinline__ T from_float(float v);
template <>
__device__ __forceinline__ float from_float<float>(float v) {
  return v;
}
template <>
__device__ __forceinline__ __nv_bfloat16 from_float<__nv_bfloat16>(float v) {
  return __float2bfloat16_rn(v);
}

// CTA = 32 rows x 128 columns; 256 threads (8 warps x 4 rows each, lane = column group)
template <typename TIn, typename TOut>
__global__ void __launch_bounds__(256) pack_kernel(const __grid_constant__ PackArgs args, int64_t rows, int D) {
  const PackOperand& a = args.op[blockIdx.y];
  __shared__ float tile[32][129];
  const int64_t r0 = (int64_t)blockIdx.x * 32;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  TOut* out = static_cast<TOut*>(a.out);
  TOut* outT = static_cast<TOut*>(a.outT);
  for (int c0 = 0; c0 < D; c0 += 128) {
#pragma unroll
    for (int rr = 0; rr < 4; ++rr) {
      const int lr = w * 4 + rr;
      const int64_t r = r0 + lr;
      if (r < rows) {
        int k = 0;
        while (k + 1 < a.n && r >= a.row_begin[k + 1]) ++k;
        const TIn* src = static_cast<const TIn*>(a.ptr[k]) + (r - a.row_begin[k]) * a.row_stride[k];
#pragma unroll
        for (int cc = 0; cc < 4; ++cc) {
          const int c = c0 + cc * 32 + lane;
          if (c < D) {
            const float v = (float)src[(int64_t)c * a.col_stride[k]];
            out[r * D + c] = from_float<TOut>(v);
            tile[lr][cc * 32 + lane] = v;
          }
        }
      }
    }
    if (outT) {
      __syncthreads();
      // 128 columns x 32 rows: warp w writes columns w, w+8, ...; lane = row
#pragma unroll 4
      for (int cc = w; cc < 128; cc += 8) {
        const int c = c0 + cc;
        const int64_t r = r0 + lane;
        if (c < D && r < rows) outT[(int64_t)c * a.ldT + r] = from_float<TOut>(tile[lane][cc]);
      }
      __syncthreads();
    }
  }
}

// Fast path (the shapes the training step produces): fp32 rows with unit column stride, 16-byte aligned, D % 4 == 0 and no
// transposed copy.  One float4 per thread -> four packed outputs; at B = 65 536 the generic kernel above took 31 us per
// operand for 50 MB of traffic, 4x the HBM time.
template <typename TOut>
__global__ void __launch_bounds__(256) pack_rows_vec_kernel(const __grid_constant__ PackArgs args, int64_t rows, int D) {
  const PackOperand& a = args.op[blockIdx.y];
  const int q = D >> 2;                                   // float4 items per row
  const int64_t total = rows * q;
  for (int64_t it = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; it < total; it += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = it / q;
    const int c = (int)(it - r * q) << 2;
    int k = 0;
    while (k + 1 < a.n && r >= a.row_begin[k + 1]) ++k;
    const float* src = static_cast<const float*>(a.ptr[k]) + (r - a.row_begin[k]) * a.row_stride[k] + c;
    const float4 v = *reinterpret_cast<const float4*>(src);
    if constexpr (sizeof(TOut) == 4) {
      *reinterpret_cast<float4*>(static_cast<float*>(a.out) + r * D + c) = v;
    } else {
      __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
      uint2 o;
      o.x = *reinterpret_cast<uint32_t*>(&lo);
      o.y = *reinterpret_cast<uint32_t*>(&hi);
      *reinterpret_cast<uint2*>(static_cast<__nv_bfloat16*>(a.out) + r * D + c) = o;
    }
  }
}

static bool pack_vec_ok(const PackArgs& a, int n_ops, int D) {
  if (D % 4) return false;
  for (int i = 0; i < n_ops; ++i) {
    const PackOperand& o = a.op[i];
    if (o.outT || (reinterpret_cast<uintptr_t>(o.out) & 15)) return false;
    for (int k = 0; k < o.n; ++k)
      if (o.col_stride[k] != 1 || (o.row_stride[k] & 3) || (reinterpret_cast<uintptr_t>(o.ptr[k]) & 15)) return false;
  }
  return true;
}

template <typename TIn, typename TOut>
static int launch_pack(const PackArgs& a, int n_ops, int64_t rows, int D, cudaStream_t st) {
  if constexpr (sizeof(TIn) == 4) {
    if (pack_vec_ok(a, n_ops, D)) {
      const int64_t blocks = std::min<int64_t>(ceil_div(rows * (D / 4), 256), (int64_t)device_sm_count() * 8);
      pack_rows_vec_kernel<TOut><<<dim3((unsigned)blocks, (unsigned)n_ops), 256, 0, st>>>(a, rows, D);
      PCLIP_CHECK_LAUNCH();
      return PCLIP_OK;
    }
  }
  dim3 grid((unsigned)ceil_div(rows, 32), (unsigned)n_ops);
  pack_kernel<TIn, TOut><<<grid, 256, 0, st>>>(a, rows, D);
  PCLIP_CHECK_LAUNCH();
  return PCLIP_OK;
}

static int fill_operand(PackOperand& o, const pclip_chunk_t* chunks, int n, int64_t rows, void* out, void* outT,
                        int64_t ldT) {
  PCLIP_REQUIRE(chunks && out && n > 0, PCLIP_ERR_INVALID, "pack: null argument");
  PCLIP_REQUIRE(n <= kMaxChunks, PCLIP_ERR_UNSUPPORTED, "pack: at most %d chunks (got %d)", kMaxChunks, n);
  PCLIP_REQUIRE(!outT || ldT >= rows, PCLIP_ERR_INVALID, "pack: ldT < rows");
  o.n = n;
  o.out = out;
  o.outT = outT;
  o.ldT = ldT;
  int64_t r = 0;
  for (int i = 0; i < n; ++i) {
    PCLIP_REQUIRE(chunks[i].ptr && chunks[i].rows >= 0, PCLIP_ERR_INVALID, "pack: bad chunk %d", i);
    o.ptr[i] = chunks[i].ptr;
    o.row_begin[i] = r;
    o.row_stride[i] = chunks[i].row_stride;
    o.col_stride[i] = chunks[i].col_stride;
    r += chunks[i].rows;
  }
  o.row_begin[n] = r;
  PCLIP_REQUIRE(r == rows, PCLIP_ERR_INVALID, "pack: chunk rows sum to %lld, expected %lld", (long long)r,
                (long long)rows);
  return PCLIP_OK;
}

// packs one or two operands of identical shape with a single launch
int pack_operands(const pclip_chunk_t* x, int nx, const pclip_chunk_t* y, int ny, int in_dtype, int64_t rows, int D,
                  int out_dtype, void* outx, void* outxT, void* outy, void* outyT, int64_t ldT, cudaStream_t st) {
  PackArgs a;
  PCLIP_TRY(fill_operand(a.op[0], x, nx, rows, outx, outxT, ldT));
  int n_ops = 1;
  if (y) {
    PCLIP_TRY(fill_operand(a.op[1], y, ny, rows, outy, outyT, ldT));
    n_ops = 2;
  }
  if (in_dtype == PCLIP_F32 && out_dtype == PCLIP_F32) return launch_pack<float, float>(a, n_ops, rows, D, st);
  if (in_dtype == PCLIP_F32 && out_dtype == PCLIP_BF16) return launch_pack<float, __nv_bfloat16>(a, n_ops, rows, D, st);
  if (in_dtype == PCLIP_BF16 && out_dtype == PCLIP_F32) return launch_pack<__nv_bfloat16, float>(a, n_ops, rows, D, st);
  if (in_dtype == PCLIP_BF16 && out_dtype == PCLIP_BF16)
    return launch_pack<__nv_bfloat16, __nv_bfloat16>(a, n_ops, rows, D, st);
  return fail(PCLIP_ERR_UNSUPPORTED, "pack: dtype %d -> %d", in_dtype, out_dtype);
}

// ---------------------------------------------------------------------------------------------
// normalise (+ exp(tau/2)) forward / backward.  ref: modules/models.py:50-51 and :238-239
// ---------------------------------------------------------------------------------------------
__global__ void normalize_fwd_kernel(const float* __restrict__ e, int64_t rows, int D, const float* __restrict__ tau,
                                     float eps, float* __restrict__ out, float* __restrict__ inv_norm) {
  const int lane = threadIdx.x & 31;
  const int64_t row = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  if (row >= rows) return;
  const float* src = e + row * D;
  float ss = 0.f;
  for (int c = lane; c < D; c += 32) ss = fmaf(src[c], src[c], ss);
  ss = warp_sum(ss);
  const float nrm = sqrtf(ss);
  const float inv = 1.0f / fmaxf(nrm, eps);  // eps == 0: plain division (inf/nan on a zero row, as the reference)
  const float s = tau ? expf(0.5f * tau[0]) : 1.0f;
  for (int c = lane; c < D; c += 32) out[row * D + c] = src[c] * inv * s;
  if (lane == 0 && inv_norm) inv_norm[row] = (nrm > eps || eps == 0.f) ? inv : -inv;  // sign flags the clamped case
}

// Grid-wide sum with a FIXED summation order (run-to-run reproducible, unlike float atomicAdd): every block parks its
// partial, the block that takes the last ticket adds all partials in index order.  `scratch` = [gridDim.x] floats followed
// by one ticket counter that the host zeroes before the launch (atomicInc wraps it back to 0).  v is read from thread 0.
template <bool kAccumulate = true>
__device__ __forceinline__ void ordered_grid_sum(float v, float* scratch, float scale, float* out) {
  __shared__ bool last;
  unsigned* ticket = reinterpret_cast<unsigned*>(scratch + gridDim.x);
  if (threadIdx.x == 0) {
    scratch[blockIdx.x] = v;
    __threadfence();
    last = atomicInc(ticket, gridDim.x - 1) == gridDim.x - 1;
  }
  __syncthreads();
  if (last && threadIdx.x < 32) {
    __threadfence();
    float s = 0.f;
    for (unsigned i = threadIdx.x; i < gridDim.x; i += 32) s += *reinterpret_cast<volatile float*>(scratch + i);
    s = warp_sum(s);
    if (threadIdx.x == 0) *out = kAccumulate ? *out + s * scale : s * scale;
  }
}

// de = (s * inv) * (g - nhat <g, nhat>)  (projection dropped when the norm was clamped to eps);
// dtau += 0.5 * sum(g * out)
__global__ void normalize_bwd_kernel(const float* __restrict__ g, const float* __restrict__ out,
                                     const float* __restrict__ inv_norm, int64_t rows, int D,
                                     const float* __restrict__ tau, float* __restrict__ de, float* __restrict__ dtau,
                                     float* __restrict__ scratch) {
  const int lane = threadIdx.x & 31;
  const int64_t row = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  float go = 0.f;
  if (row < rows) {
    const float s = tau ? expf(0.5f * tau[0]) : 1.0f;
    const float invs = 1.0f / s;
    const float invn = inv_norm[row];
    const bool clamped = invn < 0.f;
    const float inv = fabsf(invn);
    for (int c = lane; c < D; c += 32) go = fmaf(g[row * D + c], out[row * D + c], go);
    go = warp_sum(go);  // <g, out> = s <g, nhat>
    const float proj = clamped ? 0.f : go * invs;
    for (int c = lane; c < D; c += 32) {
      const float nhat = out[row * D + c] * invs;
      de[row * D + c] = s * inv * (g[row * D + c] - nhat * proj);
    }
  }
  if (dtau) {
    __shared__ float red[32];
    const int w = threadIdx.x >> 5;
    if (lane == 0) red[w] = row < rows ? go : 0.f;
    __syncthreads();
    float v = 0.f;
    if (w == 0) {
      v = lane < (int)(blockDim.x >> 5) ? red[lane] : 0.f;
      v = warp_sum(v);
    }
    ordered_grid_sum(v, scratch, 0.5f, dtau);
  }
}

// ---------------------------------------------------------------------------------------------
// CLIP finalise.  ref: modules/training_utils.py:87-89
// ---------------------------------------------------------------------------------------------
// Per-row outputs on a full grid; the loss is then summed by ONE block in a fixed order (deterministic) from the
// log-sums just written.  (All of it in a single block took 45 us at M = 65 536: one SM's issue rate.)
__global__ void clip_finalize_rows_kernel(const float* __restrict__ rowsum, const float* __restrict__ colsum, int64_t M,
                                          float* __restrict__ row_lse, float* __restrict__ col_lse,
                                          float* __restrict__ rinv, float* __restrict__ cinv) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= M) return;
  const float r = rowsum[i], c = colsum[i];
  row_lse[i] = logf(r);
  col_lse[i] = logf(c);
  if (rinv) rinv[i] = 1.0f / r;
  if (cinv) cinv[i] = 1.0f / c;
}

__global__ void __launch_bounds__(1024) clip_loss_sum_kernel(const float* __restrict__ row_lse,
                                                             const float* __restrict__ col_lse,
                                                             const float* __restrict__ diag, int64_t M, float scale,
                                                             float* __restrict__ loss) {
  __shared__ float red[32];
  float acc = 0.f;
  constexpr int kBatch = 8;
  for (int64_t i0 = threadIdx.x; i0 < M; i0 += (int64_t)blockDim.x * kBatch) {
    float lr[kBatch], lc[kBatch], d[kBatch];
#pragma unroll
    for (int u = 0; u < kBatch; ++u) {
      const int64_t i = i0 + (int64_t)u * blockDim.x;
      const bool ok = i < M;
      lr[u] = ok ? row_lse[i] : 0.f;
      lc[u] = ok ? col_lse[i] : 0.f;
      d[u] = ok ? diag[i] : 0.f;
    }
#pragma unroll
    for (int u = 0; u < kBatch; ++u) acc += (lr[u] - d[u]) + (lc[u] - d[u]);   // ascending rows per thread
  }
  acc = warp_sum(acc);
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (lane == 0) red[w] = acc;
  __syncthreads();
  if (w == 0) {
    float v = lane < (int)(blockDim.x >> 5) ? red[lane] : 0.f;
    v = warp_sum(v);
    if (lane == 0) loss[0] = v * scale;
  }
}

// Row shard, after the all-gather of every rank's record [C partial (Nb) | R (Ma) | S_ii (Ma)]: reduce the column
// partials in rank order, take R and S_ii by global row, finalise all Nb rows and sum the loss terms in a fixed order
// (per block, then blocks in index order by the last block) -- one launch instead of a reduction, two strided copies, the
// per-row kernel and a single-block sum over 3 x Nb floats (12-24 us at Nb = 65 536).
__global__ void __launch_bounds__(256) clip_finalize_sharded_kernel(const float* __restrict__ stats, int P, int64_t Nb,
                                                                    int64_t Ma, float scale, float* __restrict__ row_lse,
                                                                    float* __restrict__ col_lse, float* __restrict__ rinv,
                                                                    float* __restrict__ cinv, float* __restrict__ scratch,
                                                                    float* __restrict__ loss) {
  __shared__ float red[8];
  const int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  float term = 0.f;
  if (j < Nb) {
    const int64_t rec = Nb + 2 * Ma;
    float c = 0.f;
    for (int p = 0; p < P; ++p) c += stats[p * rec + j];
    const int64_t p = j / Ma, i = j - p * Ma;
    const float r = stats[p * rec + Nb + i], d = stats[p * rec + Nb + Ma + i];
    const float lr = logf(r), lc = logf(c);
    row_lse[j] = lr;
    col_lse[j] = lc;
    rinv[j] = 1.0f / r;
    cinv[j] = 1.0f / c;
    term = (lr - d) + (lc - d);
  }
  term = warp_sum(term);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = term;
  __syncthreads();
  float v = 0.f;
  if (threadIdx.x == 0)
    for (int w = 0; w < 8; ++w) v += red[w];
  ordered_grid_sum<false>(v, scratch, scale, loss);
}

// single-block variant for callers that do not want the log-sums stored
__global__ void __launch_bounds__(1024) clip_finalize_kernel(const float* __restrict__ rowsum, const float* __restrict__ colsum,
                                     const float* __restrict__ diag, int64_t M, float scale,
                                     float* __restrict__ row_lse, float* __restrict__ col_lse,
                                     float* __restrict__ rinv, float* __restrict__ cinv, float* __restrict__ loss) {
  __shared__ float red[32];
  float acc = 0.f;
  for (int64_t i = threadIdx.x; i < M; i += blockDim.x) {
    const float r = rowsum[i], c = colsum[i];
    const float lr = logf(r), lc = logf(c);
    if (row_lse) row_lse[i] = lr;
    if (col_lse) col_lse[i] = lc;
    if (rinv) rinv[i] = 1.0f / r;
    if (cinv) cinv[i] = 1.0f / c;
    acc += (lr - diag[i]) + (lc - diag[i]);
  }
  acc = warp_sum(acc);
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (lane == 0) red[w] = acc;
  __syncthreads();
  if (w == 0) {
    float v = lane < (int)(blockDim.x >> 5) ? red[lane] : 0.f;
    v = warp_sum(v);
    if (lane == 0) loss[0] = v * scale;
  }
}

__global__ void scale_by_scalar_kernel(const float* __restrict__ in, const float* __restrict__ g, int64_t n,
                                       float* __restrict__ out) {
  const float s = g[0];
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    out[i] = in[i] * s;
}

// ---------------------------------------------------------------------------------------------
// row-wise InfoNCE on [B, Bc] logits.  ref: modules/training_utils_filip.py:15-23
// (max-subtracted LSE: equal to the reference's naive exp wherever that is finite)
// ---------------------------------------------------------------------------------------------
__global__ void rowwise_infonce_kernel(const float* __restrict__ logits, int64_t B, int64_t Bc, int64_t diag_offset,
                                       float scale, float* __restrict__ loss, float* __restrict__ dlogits,
                                       float* __restrict__ scratch) {
  const int lane = threadIdx.x & 31;
  const int64_t row = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  float term = 0.f;
  if (row < B) {
    const float* L = logits + row * Bc;
    float m = -FLT_MAX;
    for (int64_t c = lane; c < Bc; c += 32) m = fmaxf(m, L[c]);
    m = warp_max(m);
    float s = 0.f;
    for (int64_t c = lane; c < Bc; c += 32) s += expf(L[c] - m);
    s = warp_sum(s);
    const float lse = m + logf(s);
    const int64_t dcol = diag_offset + row;
    term = lse - L[dcol];
    if (dlogits) {
      const float inv = 1.0f / s;
      for (int64_t c = lane; c < Bc; c += 32)
        dlogits[row * Bc + c] = scale * (expf(L[c] - m) * inv - (c == dcol ? 1.f : 0.f));
    }
  }
  __shared__ float red[32];
  const int w = threadIdx.x >> 5;
  if (lane == 0) red[w] = row < B ? term : 0.f;
  __syncthreads();
  float v = 0.f;
  if (w == 0) {
    v = lane < (int)(blockDim.x >> 5) ? red[lane] : 0.f;
    v = warp_sum(v);
  }
  ordered_grid_sum(v, scratch, scale, loss);
}

// ---------------------------------------------------------------------------------------------
// Scratch of pclip_rowwise_infonce_noalloc: the per-block partials + ticket of the ordered grid sum for up to
// kNoAllocMaxRows rows, resident in every GPU that loads the library (16 KB + 256 B).  In HOST code `g_noalloc_scratch`
// is only a stand-in for the symbol: the device address comes from cudaGetSymbolAddress, once per device.
// ---------------------------------------------------------------------------------------------
constexpr int64_t kNoAllocMaxRows = 32768;
constexpr int kNoAllocFloats = (int)(kNoAllocMaxRows / 8) + 64;      // one partial per block of 8 rows + the ticket
static __device__ __align__(256) float g_noalloc_scratch[kNoAllocFloats];

constexpr int kNoAllocMaxDevices = 64;
static float* g_noalloc_ptr[kNoAllocMaxDevices];                    // device address of the array, per device ordinal
static std::mutex g_noalloc_mu;

static int noalloc_scratch(float** out) {
  int dev = -1;
  PCLIP_CHECK_CUDA(cudaGetDevice(&dev));
  PCLIP_REQUIRE(dev >= 0 && dev < kNoAllocMaxDevices, PCLIP_ERR_UNSUPPORTED, "noalloc: device ordinal %d not supported", dev);
  std::lock_guard<std::mutex> lock(g_noalloc_mu);
  if (!g_noalloc_ptr[dev]) {
    void* p = nullptr;
    PCLIP_CHECK_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));   // of the CURRENT device's copy of the array
    PCLIP_REQUIRE(p != nullptr, PCLIP_ERR_CUDA, "noalloc: cudaGetSymbolAddress returned a null address");
    g_noalloc_ptr[dev] = static_cast<float*>(p);
  }
  *out = g_noalloc_ptr[dev];
  return PCLIP_OK;
}

}  // namespace pclip

using namespace pclip;

extern "C" {

int pclip_pack(const pclip_chunk_t* chunks, int32_t n, int32_t in_dtype, int64_t rows, int32_t D, int32_t out_dtype,
               void* out, void* outT, int64_t ldT, pclip_stream_t stream) {
  PCLIP_TRY(require_device());
  PCLIP_REQUIRE(rows > 0 && D > 0, PCLIP_ERR_INVALID, "pack: rows and D must be positive");
  return pack_operands(chunks, n, nullptr, 0, in_dtype, rows, D, out_dtype, out, outT, nullptr, nullptr, ldT,
                       static_cast<cudaStream_t>(stream));
}

int pclip_pack_pair(const pclip_chunk_t* x_chunks, int32_t nx, const pclip_chunk_t* y_chunks, int32_t ny, int32_t in_dtype,
                    int64_t rows, int32_t D, int32_t out_dtype, void* out_x, void* out_y, pclip_stream_t stream) {
  PCLIP_TRY(require_device());
  PCLIP_REQUIRE(rows > 0 && D > 0 && y_chunks, PCLIP_ERR_INVALID, "pack_pair: bad argument");
  return pack_operands(x_chunks, nx, y_chunks, ny, in_dtype, rows, D, out_dtype, out_x, nullptr, out_y, nullptr, 0,
                       static_cast<cudaStream_t>(stream));
}

int pclip_transpose_bf16(const void* in, int64_t rows, int32_t D, void* outT, int64_t ldT, pclip_stream_t stream) {
  PCLIP_TRY(require_device());
  PCLIP_REQUIRE(in && outT && rows > 0 && D > 0 && ldT >= rows, PCLIP_ERR_INVALID, "transpose: bad argument");
  pclip_chunk_t c = {in, rows, D, 1};
  // a pack whose packed output goes back onto the input (same values) and whose transposed copy is outT
  return pack_operands(&c, 1, nullptr, 0, PCLIP_BF16, rows, D, PCLIP_BF16, const_cast<void*>(in), outT, nullptr,
                       nullptr, ldT, static_cast<cudaStream_t>(stream));
}

int pclip_normalize_scale_fwd(const float* e, int64_t rows, int32_t D, const float* tau, float eps, float* out,
                              float* inv_norm, pclip_stream_t stream) {
  PCLIP_TRY(require_device());
  PCLIP_REQUIRE(e && out && rows >= 0 && D > 0, PCLIP_ERR_INVALID, "normalize_fwd: bad argument");
  if (rows == 0) return PCLIP_OK;
  const int threads = 256;
  normalize_fwd_kernel<<<(unsigned)ceil_div(rows * 32, threads), threads, 0, static_cast<cudaStream_t>(stream)>>>(
      e, rows, D, tau, eps, out, inv_norm);
  PCLIP_CHECK_LAUNCH();
  return PCLIP_OK;
}

// scratch of the ordered grid sums: one float per block of 8 rows + the ticket counter
size_t pclip_rowsum_workspace_bytes(int64_t rows) { return align_up((size_t)(ceil_div(rows, 8) + 1) * sizeof(float), 1024); }

static int zero_ticket(void* workspace, size_t workspace_bytes, int64_t rows, cudaStream_t st, float** scratch) {
  PCLIP_REQUIRE(workspace && (reinterpret_cast<uintptr_t>(workspace) & 1023) == 0 &&
                    workspace_bytes >= pclip_rowsum_workspace_bytes(rows),
                PCLIP_ERR_WORKSPACE, "workspace missing, misaligned or too small (pclip_rowsum_workspace_bytes)");
  *scratch = static_cast<float*>(workspace);
  PCLIP_CHECK_CUDA(cudaMemsetAsync(*scratch + ceil_div(rows, 8), 0, sizeof(unsigned), st));
  return PCLIP_OK;
}

int pclip_normalize_scale_bwd(const float* grad_out, const float* out, const float* inv_norm, int64_t rows, int32_t D,
                              const float* tau, float* de, float* dtau, void* workspace, size_t workspace_bytes,
                              pclip_stream_t stream) {
  PCLIP_TRY(require_device());
  PCLIP_REQUIRE(grad_out && out && inv_norm && de && rows >= 0 && D > 0, PCLIP_ERR_INVALID,
                "normalize_bwd: bad argument");
  if (rows == 0) return PCLIP_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  float* scratch = nullptr;
  if (dtau) PCLIP_TRY(zero_ticket(workspace, workspace_bytes, rows, st, &scratch));
  const int threads = 256;
  normalize_bwd_kernel<<<(unsigned)ceil_div(rows * 32, threads), threads, 0, st>>>(grad_out, out, inv_norm, rows, D, tau,
                                                                                  de, dtau, scratch);
  PCLIP_CHECK_LAUNCH();
  return PCLIP_OK;
}

int pclip_clip_finalize(const float* rowsum, const float* colsum, const float* diag, int64_t M, float scale,
                        float* row_lse, float* col_lse, float* rinv, float* cinv, float* loss, pclip_stream_t stream) {
  PCLIP_TRY(require_device());
  PCLIP_REQUIRE(rowsum && colsum && diag && loss && M > 0, PCLIP_ERR_INVALID, "finalize: bad argument");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (row_lse && col_lse && M >= 4096) {
    clip_finalize_rows_kernel<<<(unsigned)ceil_div(M, 256), 256, 0, st>>>(rowsum, colsum, M, row_lse, col_lse, rinv, cinv);
    PCLIP_CHECK_LAUNCH();
    clip_loss_sum_kernel<<<1, 1024, 0, st>>>(row_lse, col_lse, diag, M, scale, loss);
  } else {
    clip_finalize_kernel<<<1, 1024, 0, st>>>(rowsum, colsum, diag, M, scale, row_lse, col_lse, rinv, cinv, loss);
  }
  PCLIP_CHECK_LAUNCH();
  return PCLIP_OK;
}

size_t pclip_clip_finalize_sharded_workspace_bytes(int64_t Nb) {
  return align_up((size_t)(ceil_div(Nb, 256) + 1) * sizeof(float), 1024);
}

int pclip_clip_finalize_sharded(const float* stats_all, int32_t P, int64_t Nb, int64_t Ma, float scale, float* row_lse,
                                float* col_lse, float* rinv, float* cinv, float* loss, void* workspace,
                                size_t workspace_bytes, pclip_stream_t stream) {
  PCLIP_TRY(require_device());
  PCLIP_REQUIRE(stats_all && row_lse && col_lse && rinv && cinv && loss, PCLIP_ERR_INVALID,
                "finalize_sharded: null argument");
  PCLIP_REQUIRE(P >= 1 && Ma > 0 && Nb == (int64_t)P * Ma, PCLIP_ERR_INVALID,
                "finalize_sharded: Nb must be P * Ma (got P=%d Ma=%lld Nb=%lld)", P, (long long)Ma, (long long)Nb);
  PCLIP_REQUIRE(workspace && (reinterpret_cast<uintptr_t>(workspace) & 1023) == 0 &&
                    workspace_bytes >= pclip_clip_finalize_sharded_workspace_bytes(Nb),
                PCLIP_ERR_WORKSPACE, "finalize_sharded: workspace missing, misaligned or too small");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  float* scratch = static_cast<float*>(workspace);
  const unsigned blocks = (unsigned)ceil_div(Nb, 256);
  PCLIP_CHECK_CUDA(cudaMemsetAsync(scratch + blocks, 0, sizeof(unsigned), st));   // the ticket of the ordered sum
  clip_finalize_sharded_kernel<<<blocks, 256, 0, st>>>(stats_all, P, Nb, Ma, scale, row_lse, col_lse, rinv, cinv, scratch,
                                                        loss);
  PCLIP_CHECK_LAUNCH();
  return PCLIP_OK;
}

int pclip_scale_by_scalar(const float* in, const float* g, int64_t n, float* out, pclip_stream_t stream) {
  PCLIP_TRY(require_device());
  PCLIP_REQUIRE(in && g && out && n >= 0, PCLIP_ERR_INVALID, "scale_by_scalar: bad argument");
  if (n == 0) return PCLIP_OK;
  const int blocks = (int)std::min<int64_t>(ceil_div(n, 256), (int64_t)device_sm_count() * 16);
  scale_by_scalar_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(in, g, n, out);
  PCLIP_CHECK_LAUNCH();
  return PCLIP_OK;
}

int pclip_rowwise_infonce(const float* logits, int64_t B, int64_t Bc, int64_t diag_offset, float scale, float* loss,
                          float* dlogits, void* workspace, size_t workspace_bytes, pclip_stream_t stream) {
  PCLIP_TRY(require_device());
  PCLIP_REQUIRE(logits && loss && B > 0 && Bc > 0 && diag_offset >= 0 && diag_offset + B <= Bc, PCLIP_ERR_INVALID,
                "rowwise_infonce: bad argument");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  float* scratch = nullptr;
  PCLIP_TRY(zero_ticket(workspace, workspace_bytes, B, st, &scratch));
  const int threads = 256;
  rowwise_infonce_kernel<<<(unsigned)ceil_div(B * 32, threads), threads, 0, st>>>(logits, B, Bc, diag_offset, scale, loss,
                                                                                dlogits, scratch);
  PCLIP_CHECK_LAUNCH();
  return PCLIP_OK;
}

int64_t pclip_rowwise_infonce_noalloc_max_rows(void) { return kNoAllocMaxRows; }

int pclip_rowwise_infonce_noalloc(const float* logits, int64_t B, int64_t Bc, int64_t diag_offset, float scale, float* loss,
                                  float* dlogits, pclip_stream_t stream) {
  PCLIP_TRY(require_device());
  PCLIP_REQUIRE(logits && loss && B > 0 && Bc > 0 && diag_offset >= 0 && diag_offset + B <= Bc, PCLIP_ERR_INVALID,
                "rowwise_infonce_noalloc: bad argument");
  PCLIP_REQUIRE(B <= kNoAllocMaxRows, PCLIP_ERR_UNSUPPORTED,
                "rowwise_infonce_noalloc: B = %lld exceeds the static scratch (at most %lld rows); use "
                "pclip_rowwise_infonce with a caller-owned workspace", (long long)B, (long long)kNoAllocMaxRows);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  float* scratch = nullptr;
  PCLIP_TRY(noalloc_scratch(&scratch));
  // the array keeps the previous call's contents: the partials are overwritten before they are read, the ticket
  // must start at zero
  PCLIP_CHECK_CUDA(cudaMemsetAsync(scratch + ceil_div(B, 8), 0, sizeof(unsigned), st));
  const int threads = 256;
  rowwise_infonce_kernel<<<(unsigned)ceil_div(B * 32, threads), threads, 0, st>>>(logits, B, Bc, diag_offset, scale, loss,
                                                                                dlogits, scratch);
  PCLIP_CHECK_LAUNCH();
  return PCLIP_OK;
}

}  // extern "C"
