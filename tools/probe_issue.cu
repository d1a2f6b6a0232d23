// Hardware probe (debug aid): tcgen05.mma cost per instruction when the issuing code is as lean as possible
// (descriptors precomputed, no runtime branching), for `if (lane == 0)` issue versus warp-uniform code + elect.sync.
// build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I protein-clip_b200/csrc -I include
//        tools/probe_issue.cu -o build/probe_issue -lcuda
#include <stdio.h>
#include <stdlib.h>

#include "ptx.cuh"
using namespace pclip;

// MODE 0: A K-major, B K-major;  1: A K-major, B MN-major;  2: A MN-major, B K-major (N <= 64: B inside one panel row set)
template <int N, int MODE, int UNIFORM, int EXTRA = 0>   // EXTRA 1: tcgen05.fence::after_thread_sync per group; 2: + commit per group; 3: + mbarrier wait on the group before last
__global__ void probe(int reps, long long* cycles) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* a = smem;               // 32 KB
  uint8_t* b = smem + 32768;       // 64 KB
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 32768 * 3);
  uint32_t* slot = reinterpret_cast<uint32_t*>(bar + 2);
  for (int i = threadIdx.x; i < 32768 * 3 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (threadIdx.x == 0) {
    mbar_init(bar, 1);
    mbar_init(bar + 1, 1);
    fence_mbar_init();
  }
  fence_proxy_async_smem();
  if (threadIdx.x < 32) tmem_alloc<512>(slot);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tm = *slot;
  if (UNIFORM ? threadIdx.x < 32 : threadIdx.x == 0) {
    const uint32_t aa = smem_u32(a), ba = smem_u32(b);
    uint64_t ad[8], bd[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const uint32_t koff = (uint32_t)(k >> 2) * 16384u + (uint32_t)(k & 3) * 32u;
      if (MODE == 0) {
        ad[k] = make_smem_desc_sw128(aa + koff, 16, 1024);
        bd[k] = make_smem_desc_sw128(ba + (N > 128 ? (uint32_t)(k >> 2) * 32768u + (uint32_t)(k & 3) * 32u : koff), 16, 1024);
      } else if (MODE == 1) {
        ad[k] = make_smem_desc_sw128(aa + koff, 16, 1024);
        bd[k] = make_smem_desc_sw128(ba + (uint32_t)k * 2048u, 16384, 1024);
      } else {
        ad[k] = make_smem_desc_sw128(aa + (uint32_t)k * 2048u, 16384, 1024);
        bd[k] = make_smem_desc_sw128(ba + (uint32_t)(k & 3) * 32u, 16, 1024);
      }
    }
    constexpr uint32_t idesc = make_idesc_bf16(128, N, MODE == 2 ? 1 : 0, MODE == 1 ? 1 : 0);
    const long long t0 = clock64();
    for (int r = 0; r < reps; ++r) {
      if (EXTRA >= 3 && r >= 2) mbar_wait(bar + 1, (uint32_t)(r - 2) & 1u);
      if (EXTRA >= 1) tc_fence_after_sync();
#pragma unroll
      for (int k = 0; k < 8; ++k)
        if (!UNIFORM || elect_one()) umma_bf16_ss(tm + (uint32_t)(r & 1) * 256u, ad[k], bd[k], idesc, k > 0 ? 1u : 0u);
      if (EXTRA >= 2) umma_commit(bar + 1);
    }
    if (!UNIFORM || elect_one()) umma_commit(bar);
    mbar_wait(bar, 0);
    const long long t1 = clock64();
    if (threadIdx.x == 0) cycles[0] = t1 - t0;
  }
  tc_fence_before_sync();
  __syncthreads();
  if (threadIdx.x < 32) tmem_dealloc<512>(tm);
}

template <int N, int MODE, int UNIFORM, int EXTRA = 0>
void run(const char* name, long long* c) {
  const int smem = 32768 * 4 + 2048, reps = 2000;
  cudaFuncSetAttribute(probe<N, MODE, UNIFORM, EXTRA>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  probe<N, MODE, UNIFORM, EXTRA><<<1, 128, smem>>>(reps, c);
  cudaError_t e = cudaDeviceSynchronize();
  long long h = 0;
  cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
  printf("x%d %-8s N=%3d %-22s: %s  %6.1f clk per MMA (M128 K16)\n", EXTRA, UNIFORM ? "elect" : "lane0", N, name, cudaGetErrorString(e),
         (double)h / (reps * 8));
}

int main() {
  long long* c;
  cudaMalloc(&c, 16);
  run<16, 0, 0>("KK", c);   run<16, 0, 1>("KK", c);
  run<64, 0, 0>("KK", c);   run<64, 0, 1>("KK", c);
  run<128, 0, 0>("KK", c);  run<128, 0, 1>("KK", c);
  run<256, 0, 0>("KK", c);  run<256, 0, 1>("KK", c);
  run<128, 1, 0>("K,MN (PV-like)", c);  run<128, 1, 1>("K,MN (PV-like)", c);
  run<16, 2, 0>("MN,K (column sums)", c);  run<16, 2, 1>("MN,K (column sums)", c);
  run<128, 0, 0, 1>("KK", c);
  run<128, 0, 0, 2>("KK", c);
  run<128, 0, 0, 3>("KK", c);
  run<16, 2, 0, 3>("MN,K", c);
  return 0;
}
