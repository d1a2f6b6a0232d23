// 64 x 64 fp32 score tile on CUDA cores: one ascending-k FFMA chain per score (shared by the CLIP fp32 path,
// the score / rank / top-k kernels and the FILIP fp32 path, so that all of them produce bit-identical scores).
#pragma once

#include "common.cuh"

namespace pclip {

constexpr int kBM = 64, kBN = 64, kBK = 16, kPad = 4;

// acc[i][j] += sum_k a[row0 + ty*4+i, k] * b[col0 + tx*4+j, k]   (k ascending)
// rowmap (optional): logical row r of `a` lives at physical row rowmap[r] (Ma counts logical rows)
__device__ __forceinline__ void score_tile(const float* __restrict__ a, const float* __restrict__ b, int64_t Ma,
                                           int64_t Nb, int D, int64_t row0, int64_t col0, float (*As)[kBM + kPad],
                                           float (*Bs)[kBN + kPad], float (&acc)[4][4],
                                           const int32_t* __restrict__ rowmap = nullptr) {
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  for (int k0 = 0; k0 < D; k0 += kBK) {
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const int idx = tid + t * 256;
      const int r = idx >> 4, k = idx & 15;
      const int64_t gr = row0 + r, gc = col0 + r;
      As[k][r] = (gr < Ma && k0 + k < D) ? a[(rowmap ? (int64_t)rowmap[gr] : gr) * D + k0 + k] : 0.f;
      Bs[k][r] = (gc < Nb && k0 + k < D) ? b[gc * D + k0 + k] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < kBK; ++k) {
      const float4 av = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
      const float4 bv = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
      const float ar[4] = {av.x, av.y, av.z, av.w};
      const float br[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(ar[i], br[j], acc[i][j]);
    }
    __syncthreads();
  }
}

}  // namespace pclip
