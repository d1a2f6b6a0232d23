// FP32 CUDA-core path of the CLIP InfoNCE contractions (PCLIP_COMPUTE_F32).
//
// Every score is one ascending-d FFMA chain in fp32 and exp is expf(), so the result tracks the reference's fp32
// torch path (ref: modules/training_utils.py:85-89) to a few ulp.  Used for fp32 parity (cfg 1, evaluate()) and as
// the exact mode of any shape the tensor-core path does not take (D != 128).
//
// rowsumexp: CTA = 64 rows of `a`, loops over all 64-column tiles of `b`; row sums stay in registers, so the
//            summation order is fixed (deterministic, no atomics).
// grad:      CTA = 64 rows of `a` x 64 output columns; per column tile it recomputes the score tile, forms
//            P = exp(S) (rinv_i + cinv_j) in shared memory and accumulates P . b_tile.
#include "common.cuh"
#include "simt_tile.cuh"

namespace pclip {

struct FwdSides {
  pclip_clip_fwd_side_t s[2];
};
struct BwdSides {
  pclip_clip_bwd_side_t s[2];
};

__global__ void __launch_bounds__(256) simt_rowsumexp_kernel(FwdSides sides, int D) {
  const pclip_clip_fwd_side_t& sd = sides.s[blockIdx.z];
  const int64_t row0 = (int64_t)blockIdx.x * kBM;
  if (row0 >= sd.Ma) return;
  __shared__ __align__(16) float As[kBK][kBM + kPad];
  __shared__ __align__(16) float Bs[kBK][kBN + kPad];
  const float* a = static_cast<const float*>(sd.a);
  const float* b = static_cast<const float*>(sd.b);
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  float rsum[4] = {0.f, 0.f, 0.f, 0.f};
  float dg[4] = {0.f, 0.f, 0.f, 0.f};
  float acc[4][4];
  for (int64_t col0 = 0; col0 < sd.Nb; col0 += kBN) {
    score_tile(a, b, sd.Ma, sd.Nb, D, row0, col0, As, Bs, acc);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int64_t drow = sd.diag_offset + row0 + ty * 4 + i;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int64_t col = col0 + tx * 4 + j;
        if (col < sd.Nb) {
          rsum[i] += expf(acc[i][j]);
          if (col == drow) dg[i] = acc[i][j];
        }
      }
    }
  }
  // the 16 threads of a row group are 16 consecutive lanes
#pragma unroll
  for (int i = 0; i < 4; ++i) {
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) {
      rsum[i] += __shfl_xor_sync(0xffffffffu, rsum[i], o);
      dg[i] += __shfl_xor_sync(0xffffffffu, dg[i], o);  // exactly one lane holds a non-zero value
    }
    const int64_t row = row0 + ty * 4 + i;
    if (tx == 0 && row < sd.Ma) {
      sd.rowsum[row] = rsum[i];
      if (sd.diag) sd.diag[row] = dg[i];
    }
  }
}

__global__ void __launch_bounds__(256) simt_grad_kernel(BwdSides sides, int D, float scale, const float* gscale) {
  const pclip_clip_bwd_side_t& sd = sides.s[blockIdx.z];
  const int64_t row0 = (int64_t)blockIdx.x * kBM;
  if (row0 >= sd.Ma) return;
  const int d0 = blockIdx.y * 64;
  __shared__ __align__(16) float As[kBK][kBM + kPad];
  __shared__ __align__(16) float Bs[kBK][kBN + kPad];
  __shared__ __align__(16) float Pt[kBN][kBM + kPad];  // P transposed: [j][row]
  __shared__ __align__(16) float Bd[kBN][64 + kPad];   // b[col0 + j][d0 + c]
  const float* a = static_cast<const float*>(sd.a);
  const float* b = static_cast<const float*>(sd.b);
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  float rinv[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int64_t row = row0 + ty * 4 + i;
    rinv[i] = row < sd.Ma ? sd.rinv[row] : 0.f;
  }
  float out[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) out[i][j] = 0.f;
  float acc[4][4];
  for (int64_t col0 = 0; col0 < sd.Nb; col0 += kBN) {
    score_tile(a, b, sd.Ma, sd.Nb, D, row0, col0, As, Bs, acc);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int64_t col = col0 + tx * 4 + j;
      const float cinv = col < sd.Nb ? sd.cinv[col] : 0.f;
#pragma unroll
      for (int i = 0; i < 4; ++i)
        Pt[tx * 4 + j][ty * 4 + i] = col < sd.Nb ? expf(acc[i][j]) * (rinv[i] + cinv) : 0.f;
    }
#pragma unroll
    for (int t = 0; t < 16; ++t) {
      const int idx = tid + t * 256;
      const int j = idx >> 6, c = idx & 63;
      const int64_t gc = col0 + j;
      Bd[j][c] = (gc < sd.Nb && d0 + c < D) ? b[gc * D + d0 + c] : 0.f;
    }
    __syncthreads();
#pragma unroll 8
    for (int j = 0; j < kBN; ++j) {
      const float4 pv = *reinterpret_cast<const float4*>(&Pt[j][ty * 4]);
      const float4 bv = *reinterpret_cast<const float4*>(&Bd[j][tx * 4]);
      const float pr[4] = {pv.x, pv.y, pv.z, pv.w};
      const float br[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int c = 0; c < 4; ++c) out[i][c] = fmaf(pr[i], br[c], out[i][c]);
    }
    __syncthreads();
  }
  const float s = scale * (gscale ? gscale[0] : 1.0f);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int64_t row = row0 + ty * 4 + i;
    if (row >= sd.Ma) continue;
    const int64_t prow = sd.diag_offset + row;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const int d = d0 + tx * 4 + c;
      if (d < D) sd.da[row * D + d] = s * (out[i][c] - 2.0f * b[prow * D + d]);
    }
  }
}

int simt_rowsumexp(const pclip_clip_fwd_side_t* sides, int n_sides, int D, cudaStream_t st) {
  FwdSides fs;
  int64_t maxM = 0;
  for (int i = 0; i < n_sides; ++i) {
    fs.s[i] = sides[i];
    PCLIP_REQUIRE(sides[i].a && sides[i].b && sides[i].rowsum && sides[i].Ma > 0 && sides[i].Nb > 0,
                  PCLIP_ERR_INVALID, "rowsumexp: bad side %d", i);
    maxM = sides[i].Ma > maxM ? sides[i].Ma : maxM;
  }
  if (n_sides == 1) fs.s[1] = fs.s[0];
  dim3 grid((unsigned)ceil_div(maxM, kBM), 1, (unsigned)n_sides);
  simt_rowsumexp_kernel<<<grid, 256, 0, st>>>(fs, D);
  PCLIP_CHECK_LAUNCH();
  return PCLIP_OK;
}

int simt_grad(const pclip_clip_bwd_side_t* sides, int n_sides, int D, float scale, const float* gscale,
              cudaStream_t st) {
  BwdSides bs;
  int64_t maxM = 0;
  for (int i = 0; i < n_sides; ++i) {
    bs.s[i] = sides[i];
    PCLIP_REQUIRE(sides[i].a && sides[i].b && sides[i].rinv && sides[i].cinv && sides[i].da && sides[i].Ma > 0 &&
                      sides[i].Nb > 0,
                  PCLIP_ERR_INVALID, "grad: bad side %d", i);
    PCLIP_REQUIRE(sides[i].diag_offset >= 0 && sides[i].diag_offset + sides[i].Ma <= sides[i].Nb, PCLIP_ERR_INVALID,
                  "grad: positives of side %d fall outside b", i);
    maxM = sides[i].Ma > maxM ? sides[i].Ma : maxM;
  }
  if (n_sides == 1) bs.s[1] = bs.s[0];
  dim3 grid((unsigned)ceil_div(maxM, kBM), (unsigned)ceil_div(D, 64), (unsigned)n_sides);
  simt_grad_kernel<<<grid, 256, 0, st>>>(bs, D, scale, gscale);
  PCLIP_CHECK_LAUNCH();
  return PCLIP_OK;
}

}  // namespace pclip
