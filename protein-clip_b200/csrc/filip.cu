// FILIP token-wise late interaction: entry points, the FP32 CUDA-core path and the small tail kernels.
//   ref: modules/models.py:292-356 (_filip_similarity_score), :247-264 (_masked_mean);
//        modules/training_utils_filip.py:4-39 (row-wise InfoNCE on both similarity matrices).
//
// Both paths (fp32 here, bf16 tensor cores in filip_tc.cu) share one decomposition into two symmetric "sides":
//   side A: rows = tokens of hA[a], columns = tokens of hB[b]  -> max over s (in-thread), argA[a,b,t], partial sums
//   side B: rows = tokens of hB[b], columns = tokens of hA[a]  -> max over t (in-thread), argB[b,a,s], partial sums
// so that neither maximum needs a cross-thread reduction; a tail kernel turns the partial sums into the masked means.
// The backward is the same for both sides too: dX[x,t] = sum_y ( wX[x,y] Y[y, argX[x,y,t]] + sum_{s: argY[y,x,s]==t}
// wY[y,x] Y[y,s] ), a gather plus a scatter, with wX = gX / (max(lenX,1e-6) tau).
#include <float.h>
#include <math.h>

#include "common.cuh"
#include "simt_tile.cuh"

namespace pclip {

// ---------------------------------------------------------------------------------------------------------
// fp32 forward side: grid (ceil(Tx/64), Bx, ysplit)
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
filip_max_simt_kernel(const float* __restrict__ X, const float* __restrict__ Y, const int32_t* __restrict__ lenX,
                      const int32_t* __restrict__ lenY, int Bx, int By, int Tx, int Ty, int D, int y_per_split,
                      float* __restrict__ part /* [Bx][By][np] */, int np, int16_t* __restrict__ arg /* [Bx][By][Tx] */) {
  __shared__ __align__(16) float As[kBK][kBM + kPad];
  __shared__ __align__(16) float Bs[kBK][kBN + kPad];
  __shared__ float S[kBM][kBN + 1];
  __shared__ float red[2];
  const int bx = blockIdx.y, tile = blockIdx.x;
  const int64_t row0 = (int64_t)tile * kBM;
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int y0 = blockIdx.z * y_per_split, y1 = min(y0 + y_per_split, By);
  const int lx = min(lenX[bx], Tx);
  const float* xa = X + (int64_t)bx * Tx * D;
  float acc[4][4];
  for (int by = y0; by < y1; ++by) {
    const int ly = min(lenY[by], Ty);
    float m = -FLT_MAX;
    int am = 0;
    if (row0 < lx) {
      const float* yb = Y + (int64_t)by * Ty * D;
      for (int64_t col0 = 0; col0 < ly; col0 += kBN) {
        score_tile(xa, yb, Tx, ly, D, row0, col0, As, Bs, acc);
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) S[ty * 4 + i][tx * 4 + j] = acc[i][j];
        __syncthreads();
        if (tid < kBM) {
          const int ncol = min(kBN, ly - (int)col0);
          for (int c = 0; c < ncol; ++c) {
            const float v = S[tid][c];
            if (v > m) {  // strict: the first maximiser wins
              m = v;
              am = (int)col0 + c;
            }
          }
        }
        __syncthreads();
      }
    }
    if (tid < kBM) {
      const int64_t t = row0 + tid;
      if (t < Tx && row0 < lx) arg[((int64_t)bx * By + by) * Tx + t] = (int16_t)am;
      float v = (t < lx) ? m : 0.f;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if ((tid & 31) == 0) red[tid >> 5] = v;
    }
    __syncthreads();
    if (tid == 0) part[((int64_t)bx * By + by) * np + tile] = red[0] + red[1];
    __syncthreads();
  }
}

// sim[x,y] (written at out[x * so0 + y * so1]) = sum_p part[x][y][p] / tau / max(lenX[x], 1e-6)
__global__ void filip_finalize_kernel(const float* __restrict__ part, int np, const int32_t* __restrict__ lenX, int Bx,
                                      int By, int Tx, const float* __restrict__ tau, float* __restrict__ out,
                                      int64_t so0, int64_t so1) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= (int64_t)Bx * By) return;
  const int x = (int)(i / By), y = (int)(i % By);
  float s = 0.f;
  for (int p = 0; p < np; ++p) s += part[i * np + p];
  const float t = tau[0];
  const float cnt = fmaxf((float)min(lenX[x], Tx), 1e-6f);
  out[x * so0 + y * so1] = t > 0.f ? s / t / cnt : nanf("");
}

// ---------------------------------------------------------------------------------------------------------
// fp32 backward side: grid (ceil(Tx/32), Bx); thread == embedding dim d (strided), sole owner of dX[bx, chunk, d]
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
filip_grad_simt_kernel(const float* __restrict__ Y, const int32_t* __restrict__ lenX, const int32_t* __restrict__ lenY,
                       int Bx, int By, int Tx, int Ty, int D, const float* __restrict__ tau,
                       const float* __restrict__ gX, int64_t gx0, int64_t gx1, const float* __restrict__ gY,
                       int64_t gy0, int64_t gy1, const int16_t* __restrict__ argX /* [Bx][By][Tx] */,
                       const int16_t* __restrict__ argY /* [By][Bx][Ty] */, float* __restrict__ dX /* [Bx][Tx][D] */) {
  const int bx = blockIdx.y;
  const int t0 = blockIdx.x * 32, t1 = min(t0 + 32, Tx);
  const int lx = min(lenX[bx], Tx);
  const float inv_tau = 1.0f / tau[0];
  float* dx = dX + (int64_t)bx * Tx * D;
  for (int t = t0; t < t1; ++t)
    for (int d = threadIdx.x; d < D; d += blockDim.x) dx[(int64_t)t * D + d] = 0.f;
  if (t0 >= lx) return;
  const float cx = inv_tau / fmaxf((float)lx, 1e-6f);
  for (int by = 0; by < By; ++by) {
    const int ly = min(lenY[by], Ty);
    if (ly <= 0) continue;
    const float wx = gX[bx * gx0 + by * gx1] * cx;
    const float wy = gY[by * gy0 + bx * gy1] * inv_tau / fmaxf((float)ly, 1e-6f);
    const float* yb = Y + (int64_t)by * Ty * D;
    const int16_t* ax = argX + ((int64_t)bx * By + by) * Tx;
    const int16_t* ay = argY + ((int64_t)by * Bx + bx) * Ty;
    for (int t = t0; t < min(t1, lx); ++t) {       // gather: row t's own maximiser
      const float* src = yb + (int64_t)ax[t] * D;
      for (int d = threadIdx.x; d < D; d += blockDim.x) dx[(int64_t)t * D + d] = fmaf(wx, src[d], dx[(int64_t)t * D + d]);
    }
    for (int s = 0; s < ly; ++s) {                  // scatter: columns whose maximiser is one of this chunk's rows
      const int t = ay[s];
      if (t < t0 || t >= t1) continue;
      const float* src = yb + (int64_t)s * D;
      for (int d = threadIdx.x; d < D; d += blockDim.x) dx[(int64_t)t * D + d] = fmaf(wy, src[d], dx[(int64_t)t * D + d]);
    }
  }
}

// dtau += -(sum gA simA + sum gB simB) / tau
__global__ void filip_dtau_kernel(const float* __restrict__ gA, const float* __restrict__ simA,
                                  const float* __restrict__ gB, const float* __restrict__ simB, int64_t n,
                                  const float* __restrict__ tau, float* __restrict__ dtau) {
  __shared__ float red[32];
  float acc = 0.f;
  for (int64_t i = threadIdx.x; i < n; i += blockDim.x) acc += gA[i] * simA[i] + gB[i] * simB[i];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    float v = threadIdx.x < (blockDim.x >> 5) ? red[threadIdx.x] : 0.f;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (threadIdx.x == 0) dtau[0] += -v / tau[0];   // single block: a fixed summation order, no atomics
  }
}

int filip_finalize(const float* part, int np, const int32_t* lenX, int Bx, int By, int Tx, const float* tau, float* out,
                   int64_t so0, int64_t so1, cudaStream_t st) {
  filip_finalize_kernel<<<(unsigned)ceil_div((int64_t)Bx * By, 256), 256, 0, st>>>(part, np, lenX, Bx, By, Tx, tau, out,
                                                                                  so0, so1);
  PCLIP_CHECK_LAUNCH();
  return PCLIP_OK;
}

static int simt_filip_side_fwd(const float* X, const float* Y, const int32_t* lenX, const int32_t* lenY, int Bx, int By,
                               int Tx, int Ty, int D, const float* tau, float* part, int16_t* arg, float* out,
                               int64_t so0, int64_t so1, cudaStream_t st) {
  const int np = (int)ceil_div(Tx, kBM);
  int ysplit = (int)std::max<int64_t>(1, std::min<int64_t>(By, ceil_div(2 * device_sm_count(), (int64_t)np * Bx)));
  const int yps = (int)ceil_div(By, ysplit);
  ysplit = (int)ceil_div(By, yps);
  dim3 grid((unsigned)np, (unsigned)Bx, (unsigned)ysplit);
  filip_max_simt_kernel<<<grid, 256, 0, st>>>(X, Y, lenX, lenY, Bx, By, Tx, Ty, D, yps, part, np, arg);
  PCLIP_CHECK_LAUNCH();
  return filip_finalize(part, np, lenX, Bx, By, Tx, tau, out, so0, so1, st);
}

}  // namespace pclip

using namespace pclip;

extern "C" {

size_t pclip_filip_workspace_bytes(int64_t Ba, int64_t Bb, int64_t TA, int64_t TB, int32_t D, int32_t compute) {
  if (compute == PCLIP_COMPUTE_BF16) return tc_filip_workspace(Ba, Bb, TA, TB);
  (void)D;
  const size_t pa = (size_t)Ba * Bb * ceil_div(TA, kBM) * sizeof(float);
  const size_t pb = (size_t)Ba * Bb * ceil_div(TB, kBM) * sizeof(float);
  return align_up(pa, 1024) + align_up(pb, 1024) + 2048;
}

int pclip_filip_fwd(const float* hA, const float* hB, const int32_t* lenA, const int32_t* lenB, int64_t Ba, int64_t Bb,
                    int64_t TA, int64_t TB, int32_t D, const float* tau, int32_t compute, float* simA, float* simB,
                    int16_t* argA, int16_t* argB, void* workspace, size_t workspace_bytes, pclip_stream_t stream) {
  PCLIP_TRY(require_device());
  PCLIP_REQUIRE(hA && hB && lenA && lenB && tau && simA && simB && argA && argB, PCLIP_ERR_INVALID,
                "filip_fwd: null argument");
  PCLIP_REQUIRE(Ba > 0 && Bb > 0 && TA > 0 && TB > 0 && D > 0, PCLIP_ERR_INVALID, "filip_fwd: sizes must be positive");
  PCLIP_REQUIRE(TA <= 32767 && TB <= 32767, PCLIP_ERR_UNSUPPORTED, "filip_fwd: at most 32767 tokens per sequence");
  PCLIP_REQUIRE(workspace && (reinterpret_cast<uintptr_t>(workspace) & 1023) == 0 &&
                    workspace_bytes >= pclip_filip_workspace_bytes(Ba, Bb, TA, TB, D, compute),
                PCLIP_ERR_WORKSPACE, "filip_fwd: workspace missing, misaligned or too small");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (compute == PCLIP_COMPUTE_BF16)
    return tc_filip_fwd(hA, hB, lenA, lenB, Ba, Bb, TA, TB, D, tau, simA, simB, argA, argB, workspace, workspace_bytes, st);
  PCLIP_REQUIRE(compute == PCLIP_COMPUTE_F32, PCLIP_ERR_UNSUPPORTED, "filip_fwd: unknown compute mode %d", compute);
  Workspace w(workspace, workspace_bytes);
  float* pa = w.take<float>((size_t)Ba * Bb * ceil_div(TA, kBM));
  float* pb = w.take<float>((size_t)Ba * Bb * ceil_div(TB, kBM));
  PCLIP_TRY(simt_filip_side_fwd(hA, hB, lenA, lenB, (int)Ba, (int)Bb, (int)TA, (int)TB, D, tau, pa, argA, simA, Bb, 1, st));
  PCLIP_TRY(simt_filip_side_fwd(hB, hA, lenB, lenA, (int)Bb, (int)Ba, (int)TB, (int)TA, D, tau, pb, argB, simB, 1, Bb, st));
  return PCLIP_OK;
}

int pclip_filip_bwd(const float* hA, const float* hB, const int32_t* lenA, const int32_t* lenB, int64_t Ba, int64_t Bb,
                    int64_t TA, int64_t TB, int32_t D, const float* tau, int32_t compute, const float* simA,
                    const float* simB, const int16_t* argA, const int16_t* argB, const float* gA, const float* gB,
                    float* dhA, float* dhB, float* dtau, void* workspace, size_t workspace_bytes,
                    pclip_stream_t stream) {
  PCLIP_TRY(require_device());
  PCLIP_REQUIRE(hA && hB && lenA && lenB && tau && simA && simB && argA && argB && gA && gB && dhA && dhB,
                PCLIP_ERR_INVALID, "filip_bwd: null argument");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtau) {
    filip_dtau_kernel<<<1, 1024, 0, st>>>(gA, simA, gB, simB, Ba * Bb, tau, dtau);
    PCLIP_CHECK_LAUNCH();
  }
  if (compute == PCLIP_COMPUTE_BF16) {
    PCLIP_REQUIRE(workspace && (reinterpret_cast<uintptr_t>(workspace) & 1023) == 0 &&
                      workspace_bytes >= tc_filip_workspace(Ba, Bb, TA, TB),
                  PCLIP_ERR_WORKSPACE, "filip_bwd: workspace missing, misaligned or too small");
    return tc_filip_bwd(hA, hB, lenA, lenB, Ba, Bb, TA, TB, D, tau, argA, argB, gA, gB, dhA, dhB, workspace,
                        workspace_bytes, st);
  }
  PCLIP_REQUIRE(compute == PCLIP_COMPUTE_F32, PCLIP_ERR_UNSUPPORTED, "filip_bwd: unknown compute mode %d", compute);
  dim3 ga((unsigned)ceil_div(TA, 32), (unsigned)Ba), gb((unsigned)ceil_div(TB, 32), (unsigned)Bb);
  filip_grad_simt_kernel<<<ga, 128, 0, st>>>(hB, lenA, lenB, (int)Ba, (int)Bb, (int)TA, (int)TB, D, tau, gA, Bb, 1, gB, 1,
                                             Bb, argA, argB, dhA);
  PCLIP_CHECK_LAUNCH();
  filip_grad_simt_kernel<<<gb, 128, 0, st>>>(hA, lenB, lenA, (int)Bb, (int)Ba, (int)TB, (int)TA, D, tau, gB, 1, Bb, gA, Bb,
                                             1, argB, argA, dhB);
  PCLIP_CHECK_LAUNCH();
  return PCLIP_OK;
}

}  // extern "C"
