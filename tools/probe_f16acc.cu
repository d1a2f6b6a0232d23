// Hardware probe (debug aid, not part of the library): tcgen05.mma kind::f16 with an FP16 accumulator --
// (1) how the FP16 D tile is laid out in TMEM, (2) whether A = f16 / B = bf16 may be mixed in one instruction.
// build: nvcc -O2 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I protein-clip_b200/csrc -I include
//        tools/probe_f16acc.cu -o gpurun_out/probe_f16acc -lcuda
#include <cuda_fp16.h>
#include <stdio.h>
#include <stdlib.h>

#include "ptx.cuh"

using namespace pclip;

// mode 0: A bf16, B bf16, D f32 (control)   mode 1: A bf16, B bf16, D f16
// mode 2: A f16,  B bf16, D f32 (mixed)     mode 3: A f16, B f16, D f16
__global__ void probe(int mode, uint32_t* out /* [128 lanes][32 words] */) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* a = smem;            // [128 rows][64 x 16-bit], 128 B swizzle (only k < 16 used)
  uint8_t* b = smem + 16384;
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 32768);
  uint32_t* slot = reinterpret_cast<uint32_t*>(bar + 1);
  const bool a_f16 = mode >= 2, b_f16 = mode == 3, d_f16 = (mode == 1 || mode == 3);
  for (int idx = threadIdx.x; idx < 128 * 64; idx += blockDim.x) {
    const int r = idx / 64, k = idx % 64;
    const float av = k < 16 ? (float)(r % 8 + 1) / 8.f : 0.f;     // A[r][k]
    const float bv = k < 16 ? (float)(r % 16 + 1) / 16.f : 0.f;   // B[n = r][k]
    const uint32_t off = sw128_offset(r, k / 8) + (k % 8) * 2;
    *reinterpret_cast<uint16_t*>(a + off) = a_f16 ? __half_as_ushort(__float2half(av)) : __bfloat16_as_ushort(__float2bfloat16(av));
    *reinterpret_cast<uint16_t*>(b + off) = b_f16 ? __half_as_ushort(__float2half(bv)) : __bfloat16_as_ushort(__float2bfloat16(bv));
  }
  if (threadIdx.x == 0) {
    mbar_init(bar, 1);
    fence_mbar_init();
  }
  fence_proxy_async_smem();
  if (threadIdx.x < 32) tmem_alloc<128>(slot);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tm = *slot;
  if (threadIdx.x == 0) {
    // c_format bits [4,6): 0 = F16, 1 = F32; a_format bits [7,10), b_format bits [10,13): 0 = F16, 1 = BF16
    uint32_t idesc = ((d_f16 ? 0u : 1u) << 4) | ((a_f16 ? 0u : 1u) << 7) | ((b_f16 ? 0u : 1u) << 10) | ((128u >> 3) << 17) |
                     ((128u >> 4) << 24);
    umma_bf16_ss(tm, make_smem_desc_sw128(smem_u32(a), 16, 1024), make_smem_desc_sw128(smem_u32(b), 16, 1024), idesc, 0u);
    umma_commit(bar);
  }
  mbar_wait(bar, 0);
  tc_fence_after_sync();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint32_t r[32];
  tmem_ld_32x32b_x32(tm + ((uint32_t)(warp * 32) << 16), r);
  tmem_ld_wait();
  for (int e = 0; e < 32; ++e) out[(warp * 32 + lane) * 32 + e] = r[e];
  tc_fence_before_sync();
  __syncthreads();
  if (threadIdx.x < 32) tmem_dealloc<128>(tm);
}

int main(int argc, char** argv) {
  const int only = argc > 1 ? atoi(argv[1]) : -1;
  uint32_t* d;
  cudaMalloc(&d, 128 * 32 * 4);
  static uint32_t h[128 * 32];
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 40000);
  for (int mode = 0; mode < 4; ++mode) {
    if (only >= 0 && mode != only) continue;
    cudaMemset(d, 0xff, sizeof(h));
    probe<<<1, 128, 40000>>>(mode, d);
    cudaError_t e = cudaDeviceSynchronize();
    printf("mode %d: %s\n", mode, cudaGetErrorString(e));
    if (e != cudaSuccess) return 1;
    cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
    // expected D[r][n] = 16 * (r%8+1)/8 * (n%16+1)/16 = (r%8+1)(n%16+1)/8
    for (int lane : {0, 1, 9}) {
      printf("  lane %d words:", lane);
      for (int e = 0; e < 8; ++e) printf(" %08x", h[lane * 32 + e]);
      printf("\n    as f32:");
      for (int e = 0; e < 8; ++e) printf(" %g", *reinterpret_cast<float*>(&h[lane * 32 + e]));
      printf("\n    as f16 lo/hi:");
      for (int e = 0; e < 8; ++e) {
        __half lo = __ushort_as_half((unsigned short)(h[lane * 32 + e] & 0xffff)), hi = __ushort_as_half((unsigned short)(h[lane * 32 + e] >> 16));
        printf(" %g/%g", __half2float(lo), __half2float(hi));
      }
      printf("\n    expected cols 0..7:");
      for (int n = 0; n < 8; ++n) printf(" %g", (lane % 8 + 1) * (n % 16 + 1) / 8.0);
      printf("\n");
    }
  }
  return 0;
}
