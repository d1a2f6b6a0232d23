// Fused encoder head, forward: everything between the frozen ESM trunk and the loss,
//   projection Din -> 128, per-residue FFN ((depth-1) x [Linear, ReLU, LayerNorm] + Linear), then either
//     * the per-residue embeddings themselves                    (FILIPEncoder, ref: modules/models.py:192-196), or
//     * masked mean over the residues, pooled FFN, F.normalize, * exp(tau / 2)      (Encoder, ref: modules/models.py:46-51)
// in TWO launches instead of the ~40 of the stock nn.Sequential path.  Inference only (no dropout, nothing saved for a
// backward): the evaluation / bank-building / ranking side of the scripts; training keeps the stock autograd path.
//
// fp32 on the CUDA cores on purpose: the head is 0.1 MFLOP per residue against 300 MFLOP per residue of the ESM trunk
// in front of it, so what it costs is launches, not flops -- and fp32 keeps it within 1e-5 of the reference's fp32
// evaluation path.
//
//   head_tokens_kernel  CTA = 32 residues of one sample, 8 warps x 4 residues, lane = 4 consecutive features.  Every Linear
//                       is a k-chunked SIMT GEMM against a transposed weight chunk staged in shared memory; a residue's
//                       128 features live in ONE warp, so ReLU + LayerNorm are warp-shuffle reductions and the
//                       activations go from layer to layer through a warp-private shared-memory tile.
//   head_pooled_kernel  CTA = one sample: sum of the tile partials in tile order / clamp(count, 1), pooled FFN as
//                       warp-per-output-row mat-vecs, normalise, temperature.
#include <float.h>

#include "common.cuh"

namespace pclip {
namespace eh {

constexpr int kE = 128;          // embedding width (ref: main.py:27)
constexpr int kTT = 32;          // residues per CTA
constexpr int kKC = 32;          // k-chunk of the GEMMs
constexpr int kThreads = 256;
constexpr int kALd = kKC + 4;    // staged input chunk  [kTT][kALd]   (row stride keeps float4 alignment)
constexpr int kWLd = kE + 4;     // staged weight chunk [kKC][kWLd]   (transposed: k-major)
constexpr int kHLd = kE + 4;     // activation tile     [kTT][kHLd]
constexpr int kMaxLayers = 8;

struct Layer {
  const float* w;      // [128][K] row-major (nn.Linear.weight)
  const float* b;      // [128]
  const float* g;      // LayerNorm weight, or NULL: plain Linear (no ReLU / LayerNorm after it)
  const float* beta;   // LayerNorm bias
};

struct TokenArgs {
  const float* hidden;     // [Bs][T][Din]
  const int64_t* mask;     // [Bs][T]
  int Bs, T, Din;
  Layer proj;
  Layer ffn[kMaxLayers];
  int n_ffn;
  float ln_eps;
  float* out_tokens;       // [Bs][T][128] or NULL
  float* partial;          // [Bs][tiles][128] or NULL: masked sums of the tile's residues
};

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Ws[k][n] = W[n][k0 + k] for k < kKC (zero past K): coalesced float4 loads along k, transposing scalar stores
__device__ __forceinline__ void stage_weights(const float* __restrict__ W, int K, int k0, float* __restrict__ Ws) {
  for (int idx = threadIdx.x; idx < kE * (kKC / 4); idx += kThreads) {
    const int n = idx / (kKC / 4), k4 = (idx % (kKC / 4)) * 4;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (k0 + k4 < K) v = __ldg(reinterpret_cast<const float4*>(W + (size_t)n * K + k0 + k4));   // K % 4 == 0
    Ws[(k4 + 0) * kWLd + n] = v.x;
    Ws[(k4 + 1) * kWLd + n] = v.y;
    Ws[(k4 + 2) * kWLd + n] = v.z;
    Ws[(k4 + 3) * kWLd + n] = v.w;
  }
}

// acc[i][0..3] += sum_{k < kKC} A[4 warp + i][k] * Ws[k][4 lane .. 4 lane + 3]   (k ascending)
__device__ __forceinline__ void chunk_fma(const float* __restrict__ A, int lda, const float* __restrict__ Ws, int warp,
                                          int lane, float (&acc)[4][4]) {
#pragma unroll
  for (int k4 = 0; k4 < kKC; k4 += 4) {
    float4 a[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) a[i] = *reinterpret_cast<const float4*>(A + (warp * 4 + i) * lda + k4);
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
      const float4 w = *reinterpret_cast<const float4*>(Ws + (k4 + kk) * kWLd + lane * 4);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float av = kk == 0 ? a[i].x : kk == 1 ? a[i].y : kk == 2 ? a[i].z : a[i].w;
        acc[i][0] = fmaf(av, w.x, acc[i][0]);
        acc[i][1] = fmaf(av, w.y, acc[i][1]);
        acc[i][2] = fmaf(av, w.z, acc[i][2]);
        acc[i][3] = fmaf(av, w.w, acc[i][3]);
      }
    }
  }
}

// bias, then (for an FFN block) ReLU + LayerNorm over the 128 features of each residue -- all inside one warp
__device__ __forceinline__ void layer_epilogue(const Layer& L, float eps, int lane, float (&acc)[4][4]) {
  const float4 b = __ldg(reinterpret_cast<const float4*>(L.b) + lane);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    acc[i][0] += b.x;
    acc[i][1] += b.y;
    acc[i][2] += b.z;
    acc[i][3] += b.w;
  }
  if (L.g == nullptr) return;
  const float4 g = __ldg(reinterpret_cast<const float4*>(L.g) + lane);
  const float4 be = __ldg(reinterpret_cast<const float4*>(L.beta) + lane);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = fmaxf(acc[i][j], 0.f);
    const float mean = warp_sum((acc[i][0] + acc[i][1]) + (acc[i][2] + acc[i][3])) * (1.0f / kE);
    const float d0 = acc[i][0] - mean, d1 = acc[i][1] - mean, d2 = acc[i][2] - mean, d3 = acc[i][3] - mean;
    const float var = warp_sum((d0 * d0 + d1 * d1) + (d2 * d2 + d3 * d3)) * (1.0f / kE);    // biased, as nn.LayerNorm
    const float r = rsqrtf(var + eps);
    acc[i][0] = d0 * r * g.x + be.x;
    acc[i][1] = d1 * r * g.y + be.y;
    acc[i][2] = d2 * r * g.z + be.z;
    acc[i][3] = d3 * r * g.w + be.w;
  }
}

__global__ void __launch_bounds__(kThreads) head_tokens_kernel(const __grid_constant__ TokenArgs a) {
  __shared__ __align__(16) float As[kTT * kALd];
  __shared__ __align__(16) float Ws[kKC * kWLd];
  __shared__ __align__(16) float Hs[kTT * kHLd];
  __shared__ int s_valid;
  const int b = blockIdx.y, t0 = blockIdx.x * kTT;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tiles = gridDim.x;

  if (a.partial) {   // pooled mode: a tile without a valid residue contributes zeros and skips the arithmetic
    if (threadIdx.x == 0) s_valid = 0;
    __syncthreads();
    if (threadIdx.x < kTT && t0 + threadIdx.x < a.T && a.mask[(size_t)b * a.T + t0 + threadIdx.x] != 0) s_valid = 1;
    __syncthreads();
    if (!s_valid) {
      if (threadIdx.x < kE) a.partial[((size_t)b * tiles + blockIdx.x) * kE + threadIdx.x] = 0.f;
      return;
    }
  }

  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i) acc[i][0] = acc[i][1] = acc[i][2] = acc[i][3] = 0.f;

  // projection: the input chunk [32 residues x 32 k] is staged next to the weight chunk
  for (int k0 = 0; k0 < a.Din; k0 += kKC) {
    for (int idx = threadIdx.x; idx < kTT * (kKC / 4); idx += kThreads) {
      const int r = idx / (kKC / 4), k4 = (idx % (kKC / 4)) * 4;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (t0 + r < a.T && k0 + k4 < a.Din)
        v = __ldg(reinterpret_cast<const float4*>(a.hidden + ((size_t)b * a.T + t0 + r) * a.Din + k0 + k4));
      *reinterpret_cast<float4*>(As + r * kALd + k4) = v;
    }
    stage_weights(a.proj.w, a.Din, k0, Ws);
    __syncthreads();
    chunk_fma(As, kALd, Ws, warp, lane, acc);
    __syncthreads();
  }
  layer_epilogue(a.proj, a.ln_eps, lane, acc);

  for (int l = 0; l < a.n_ffn; ++l) {
    // this warp's four residues: registers -> its rows of the activation tile (read back only by this warp)
#pragma unroll
    for (int i = 0; i < 4; ++i)
      *reinterpret_cast<float4*>(Hs + (warp * 4 + i) * kHLd + lane * 4) =
          make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
#pragma unroll
    for (int i = 0; i < 4; ++i) acc[i][0] = acc[i][1] = acc[i][2] = acc[i][3] = 0.f;
    for (int k0 = 0; k0 < kE; k0 += kKC) {
      stage_weights(a.ffn[l].w, kE, k0, Ws);
      __syncthreads();          // also orders the Hs stores above before the first read
      chunk_fma(Hs + k0, kHLd, Ws, warp, lane, acc);
      __syncthreads();
    }
    layer_epilogue(a.ffn[l], a.ln_eps, lane, acc);
  }

  if (a.out_tokens) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int t = t0 + warp * 4 + i;
      if (t < a.T)
        *reinterpret_cast<float4*>(a.out_tokens + ((size_t)b * a.T + t) * kE + lane * 4) =
            make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
    }
  }
  if (a.partial) {
    // masked sum over the tile: residues in ascending order inside a warp, then warps in ascending order
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int t = t0 + warp * 4 + i;
      const float m = (t < a.T) ? (float)a.mask[(size_t)b * a.T + t] : 0.f;   // h * mask, as the reference (mask is 0 / 1)
      if (t < a.T) {
        s.x += acc[i][0] * m;
        s.y += acc[i][1] * m;
        s.z += acc[i][2] * m;
        s.w += acc[i][3] * m;
      }
    }
    __syncthreads();            // Hs is free: every warp is past its last chunk_fma
    *reinterpret_cast<float4*>(Hs + warp * kHLd + lane * 4) = s;
    __syncthreads();
    if (threadIdx.x < kE) {
      float v = 0.f;
#pragma unroll
      for (int w = 0; w < kThreads / 32; ++w) v += Hs[w * kHLd + threadIdx.x];
      a.partial[((size_t)b * tiles + blockIdx.x) * kE + threadIdx.x] = v;
    }
  }
}

struct PooledArgs {
  const float* partial;    // [Bs][tiles][128]
  const int64_t* mask;     // [Bs][T]
  int T, tiles;
  Layer ffn[kMaxLayers];
  int n_ffn;
  float ln_eps, norm_eps;
  const float* tau;        // [1] or NULL
  float* out;              // [Bs][128]
};

__global__ void __launch_bounds__(kE) head_pooled_kernel(const __grid_constant__ PooledArgs a) {
  __shared__ __align__(16) float xs[kE];
  __shared__ __align__(16) float ys[kE];
  __shared__ int cnt_s[kE / 32];
  const int b = blockIdx.x, n = threadIdx.x, warp = n >> 5, lane = n & 31;
  // number of valid residues (integer, exact), clamped to >= 1 (ref: modules/models.py:88)
  int c = 0;
  for (int t = n; t < a.T; t += kE) c += (int)a.mask[(size_t)b * a.T + t];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  if (lane == 0) cnt_s[warp] = c;
  float v = 0.f;
  for (int t = 0; t < a.tiles; ++t) v += a.partial[((size_t)b * a.tiles + t) * kE + n];
  __syncthreads();
  int cnt = 0;
#pragma unroll
  for (int w = 0; w < kE / 32; ++w) cnt += cnt_s[w];
  xs[n] = v / (float)max(cnt, 1);
  __syncthreads();

  for (int l = 0; l < a.n_ffn; ++l) {
    const Layer& L = a.ffn[l];
    const float4 x4 = *reinterpret_cast<const float4*>(xs + lane * 4);
    // warp w computes output rows w, w + 4, ...: one coalesced 512-byte weight row per step, all loads independent
    float dot[kE / 4];
#pragma unroll
    for (int r = 0; r < kE / 4; ++r) {
      const float4 w4 = __ldg(reinterpret_cast<const float4*>(L.w + (size_t)(r * 4 + warp) * kE) + lane);
      dot[r] = fmaf(w4.x, x4.x, fmaf(w4.y, x4.y, fmaf(w4.z, x4.z, w4.w * x4.w)));
    }
#pragma unroll
    for (int r = 0; r < kE / 4; ++r) {
      const float d = warp_sum(dot[r]);
      if (lane == 0) ys[r * 4 + warp] = d + L.b[r * 4 + warp];
    }
    __syncthreads();
    float y = ys[n];
    if (L.g) {
      y = fmaxf(y, 0.f);
      __syncthreads();
      ys[n] = y;
      __syncthreads();
      float s = (ys[lane] + ys[lane + 32]) + (ys[lane + 64] + ys[lane + 96]);
      const float mean = warp_sum(s) * (1.0f / kE);
      const float e0 = ys[lane] - mean, e1 = ys[lane + 32] - mean, e2 = ys[lane + 64] - mean, e3 = ys[lane + 96] - mean;
      const float var = warp_sum((e0 * e0 + e1 * e1) + (e2 * e2 + e3 * e3)) * (1.0f / kE);
      y = (y - mean) * rsqrtf(var + a.ln_eps) * L.g[n] + L.beta[n];
    }
    __syncthreads();
    xs[n] = y;
    __syncthreads();
  }
  // F.normalize(dim=-1, eps) * exp(tau / 2)   (ref: modules/models.py:50-51)
  const float s = (xs[lane] * xs[lane] + xs[lane + 32] * xs[lane + 32]) +
                  (xs[lane + 64] * xs[lane + 64] + xs[lane + 96] * xs[lane + 96]);
  const float nrm = sqrtf(warp_sum(s));
  const float scale = a.tau ? expf(0.5f * a.tau[0]) : 1.0f;
  a.out[(size_t)b * kE + n] = xs[n] / fmaxf(nrm, a.norm_eps) * scale;
}

}  // namespace eh
}  // namespace pclip

using namespace pclip;

extern "C" {

size_t pclip_encoder_head_workspace_bytes(int64_t Bs, int64_t T) {
  return align_up((size_t)Bs * (size_t)ceil_div(T, (int64_t)eh::kTT) * eh::kE * sizeof(float), 1024);
}

static int copy_layers(const pclip_linear_t* src, int n, eh::Layer* dst, const char* what) {
  PCLIP_REQUIRE(n >= 0 && n <= eh::kMaxLayers, PCLIP_ERR_UNSUPPORTED, "encoder_head: %s has %d layers (at most %d)", what, n,
                eh::kMaxLayers);
  PCLIP_REQUIRE(n == 0 || src, PCLIP_ERR_INVALID, "encoder_head: %s is NULL", what);
  for (int i = 0; i < n; ++i) {
    PCLIP_REQUIRE(src[i].weight && src[i].bias && ((src[i].ln_weight == nullptr) == (src[i].ln_bias == nullptr)),
                  PCLIP_ERR_INVALID, "encoder_head: %s layer %d: weight / bias missing or half a LayerNorm", what, i);
    PCLIP_REQUIRE(((reinterpret_cast<uintptr_t>(src[i].weight) | reinterpret_cast<uintptr_t>(src[i].bias) |
                    reinterpret_cast<uintptr_t>(src[i].ln_weight) | reinterpret_cast<uintptr_t>(src[i].ln_bias)) & 15) == 0,
                  PCLIP_ERR_INVALID, "encoder_head: %s layer %d: parameters must be 16-byte aligned", what, i);
    dst[i] = {src[i].weight, src[i].bias, src[i].ln_weight, src[i].ln_bias};
  }
  return PCLIP_OK;
}

int pclip_encoder_head_fwd(const float* hidden, const int64_t* mask, int64_t Bs, int64_t T, int32_t Din, int32_t E,
                           const pclip_linear_t* projection, const pclip_linear_t* token_ffn, int32_t n_token,
                           const pclip_linear_t* pooled_ffn, int32_t n_pooled, const float* tau, float ln_eps,
                           float norm_eps, float* out_tokens, float* out_pooled, void* workspace, size_t workspace_bytes,
                           pclip_stream_t stream) {
  PCLIP_TRY(require_device());
  PCLIP_REQUIRE(hidden && mask && projection && Bs >= 0 && T > 0, PCLIP_ERR_INVALID, "encoder_head: null argument");
  PCLIP_REQUIRE((out_tokens != nullptr) != (out_pooled != nullptr), PCLIP_ERR_INVALID,
                "encoder_head: exactly one of out_tokens (per-residue mode) and out_pooled (pooled mode)");
  PCLIP_REQUIRE(E == eh::kE, PCLIP_ERR_UNSUPPORTED, "encoder_head: embedding width must be %d (got %d)", eh::kE, E);
  PCLIP_REQUIRE(Din > 0 && Din % 4 == 0 && (reinterpret_cast<uintptr_t>(hidden) & 15) == 0, PCLIP_ERR_UNSUPPORTED,
                "encoder_head: Din must be a multiple of 4 and hidden 16-byte aligned (Din = %d)", Din);
  PCLIP_REQUIRE(Bs <= 65535 && T < ((int64_t)1 << 31) - eh::kTT, PCLIP_ERR_UNSUPPORTED, "encoder_head: Bs <= 65535");
  if (Bs == 0) return PCLIP_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int tiles = (int)ceil_div(T, (int64_t)eh::kTT);

  eh::TokenArgs ta{};
  ta.hidden = hidden, ta.mask = mask, ta.Bs = (int)Bs, ta.T = (int)T, ta.Din = Din, ta.ln_eps = ln_eps;
  PCLIP_TRY(copy_layers(projection, 1, &ta.proj, "projection"));
  PCLIP_REQUIRE(ta.proj.g == nullptr, PCLIP_ERR_UNSUPPORTED, "encoder_head: the projection is a plain Linear");
  PCLIP_TRY(copy_layers(token_ffn, n_token, ta.ffn, "token_ffn"));
  ta.n_ffn = n_token;
  ta.out_tokens = out_tokens;
  ta.partial = nullptr;
  if (out_pooled) {
    PCLIP_REQUIRE(workspace && (reinterpret_cast<uintptr_t>(workspace) & 1023) == 0 &&
                      workspace_bytes >= pclip_encoder_head_workspace_bytes(Bs, T),
                  PCLIP_ERR_WORKSPACE, "encoder_head: workspace missing, misaligned or too small");
    ta.partial = static_cast<float*>(workspace);
  }
  head_tokens_kernel<<<dim3((unsigned)tiles, (unsigned)Bs), eh::kThreads, 0, st>>>(ta);
  PCLIP_CHECK_LAUNCH();
  if (out_pooled) {
    eh::PooledArgs pa{};
    pa.partial = ta.partial, pa.mask = mask, pa.T = (int)T, pa.tiles = tiles, pa.ln_eps = ln_eps, pa.norm_eps = norm_eps;
    PCLIP_TRY(copy_layers(pooled_ffn, n_pooled, pa.ffn, "pooled_ffn"));
    pa.n_ffn = n_pooled;
    pa.tau = tau;
    pa.out = out_pooled;
    head_pooled_kernel<<<(unsigned)Bs, eh::kE, 0, st>>>(pa);
    PCLIP_CHECK_LAUNCH();
  }
  return PCLIP_OK;
}

}  // extern "C"
