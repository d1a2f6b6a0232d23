// Hardware probe (debug aid): tcgen05.ld throughput per SM, plain 32x32b.x32 (pack 0) vs .pack::16b (pack 1: two 16-bit accumulators
// per register) vs the same cells as two 16x256b.x4 loads (pack 2).  build: nvcc -O2 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I protein-clip_b200/csrc
//                        -I include tools/probe_ldtm.cu -o build/probe_ldtm -lcuda
#include <stdio.h>
#include <stdlib.h>

#include "ptx.cuh"
using namespace pclip;

#define LD_ARGS                                                                                                        \
  "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, %21, %22, %23, "     \
  "%24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
#define LD_OUTS                                                                                                        \
  "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),          \
      "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),           \
      "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),          \
      "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])

template <int kPack>
__global__ void probe(int iters, long long* cycles, uint32_t* sink) {
  __shared__ uint32_t slot;
  if (threadIdx.x < 32) tmem_alloc<512>(&slot);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tm = slot;
  const int warp = threadIdx.x >> 5;
  const uint32_t base = tm + ((uint32_t)((warp & 3) * 32) << 16);
  uint32_t acc = 0;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    uint32_t r[32];
    const uint32_t addr = base + (uint32_t)((it * 64 + (warp >> 2) * 32) & 255);
    if (kPack == 2) {   // the same 32 lanes x 32 columns as two 16x256b.x4 loads (filip_tc1.cu)
      asm volatile(
          "tcgen05.ld.sync.aligned.16x256b.x4.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
          : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
            "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
          : "r"(addr)
          : "memory");
      asm volatile(
          "tcgen05.ld.sync.aligned.16x256b.x4.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
          : "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
            "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
          : "r"(addr + (16u << 16))
          : "memory");
    } else if (kPack)
      asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.pack::16b.b32 " LD_ARGS : LD_OUTS : "r"(addr) : "memory");
    else
      asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 " LD_ARGS : LD_OUTS : "r"(addr) : "memory");
    tmem_ld_wait();
    acc ^= r[0] ^ r[13] ^ r[31];
  }
  __syncthreads();
  const long long t1 = clock64();
  if (threadIdx.x == 0) cycles[0] = t1 - t0;
  sink[threadIdx.x] = acc;
  tc_fence_before_sync();
  __syncthreads();
  if (threadIdx.x < 32) tmem_dealloc<512>(tm);
}

int main() {
  long long* c;
  uint32_t* s;
  cudaMalloc(&c, 8);
  cudaMalloc(&s, 4096);
  const int iters = 4096;
  for (int warps : {4, 8, 16}) {
    for (int pack = 0; pack < 3; ++pack) {
      if (pack == 2) probe<2><<<1, warps * 32>>>(iters, c, s);
      else if (pack) probe<1><<<1, warps * 32>>>(iters, c, s);
      else probe<0><<<1, warps * 32>>>(iters, c, s);
      cudaError_t e = cudaDeviceSynchronize();
      long long h = 0;
      cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
      const double regbytes = (double)warps * iters * 32 * 32 * 4;   // register bytes delivered
      printf("warps %2d pack %d: %s  %lld cycles, %.1f register-bytes/clk/SM, %.1f TMEM cells/clk/SM\n", warps, pack,
             cudaGetErrorString(e), h, regbytes / h, regbytes / 4 * (pack == 1 ? 2 : 1) / h);
    }
  }
  return 0;
}
