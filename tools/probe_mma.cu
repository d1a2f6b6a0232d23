// Hardware probe (debug aid): cycles per tcgen05.mma (cta_group::1, kind::f16, SS mode, bf16) for the shapes the CLIP
// kernels issue.  build: nvcc -O2 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I protein-clip_b200/csrc
//                        -I include tools/probe_mma.cu -o build/probe_mma -lcuda
#include <stdio.h>
#include <stdlib.h>

#include "ptx.cuh"
using namespace pclip;

// mode 0: S-like  M128 N128, A K-major, B K-major      mode 1: PV-like M128 N128, A K-major, B MN-major
// mode 2: C-like  M128 N16,  A MN-major, B K-major     mode 3: M128 N256 both K-major
// mode 4: M128 N64 both K-major                        mode 5: S-like with an idle second thread polling smem
__global__ void probe(int mode, int reps, long long* cycles, int side /* 0 none, 1 LDTM, 2 st.shared, 3 both */, int uniform) {
  __shared__ volatile int done;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* a = smem;               // 32 KB: [128][64] x 2 panels
  uint8_t* b = smem + 32768;       // 64 KB: up to N = 256
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 32768 * 3);
  uint32_t* slot = reinterpret_cast<uint32_t*>(bar + 2);
  for (int i = threadIdx.x; i < 32768 * 3 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (threadIdx.x == 0) {
    done = 0;
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
  if (uniform ? threadIdx.x < 32 : threadIdx.x == 0) {
    const uint32_t aa = smem_u32(a), ba = smem_u32(b);
    const long long t0 = clock64();
    const int mode_in = mode;
    for (int r = 0; r < reps; ++r) {
      if (mode_in == 6) mode = (r & 1) ? 2 : 0;   // alternate S-like and C-like groups
      if (mode_in == 7) mode = 0;                // S-like with a commit after every group
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const uint32_t koff = (uint32_t)(k >> 2) * 16384u + (uint32_t)(k & 3) * 32u;
        uint64_t ad, bd;
        uint32_t idesc;
        if (mode == 1) {
          ad = make_smem_desc_sw128(aa + koff, 16, 1024);
          bd = make_smem_desc_sw128(ba + (uint32_t)k * 2048u, 16384, 1024);
          idesc = make_idesc_bf16(128, 128, 0, 1);
        } else if (mode == 2) {
          ad = make_smem_desc_sw128(aa + (uint32_t)k * 2048u, 16384, 1024);
          bd = make_smem_desc_sw128(ba + (uint32_t)(k & 3) * 32u, 16, 1024);
          idesc = make_idesc_bf16(128, 16, 1, 0);
        } else {
          ad = make_smem_desc_sw128(aa + koff, 16, 1024);
          bd = make_smem_desc_sw128(ba + (mode == 3 ? (uint32_t)(k >> 2) * 32768u + (uint32_t)(k & 3) * 32u : koff), 16, 1024);
          idesc = make_idesc_bf16(128, mode == 3 ? 256 : (mode == 4 ? 64 : 128), 0, 0);
        }
        if (!uniform || elect_one()) umma_bf16_ss(tm + (uint32_t)(r & 1) * 256u, ad, bd, idesc, k > 0 ? 1u : 0u);
      }
      if (mode_in == 7 && (!uniform || elect_one())) umma_commit(bar + 1);
    }
    if (!uniform || elect_one()) umma_commit(bar);
    mbar_wait(bar, 0);
    const long long t1 = clock64();
    if (threadIdx.x == 0) { cycles[0] = t1 - t0; done = 1; }
  } else if (threadIdx.x >= 64 && side) {
    // side traffic from the other warps while the MMAs run
    const int warp = threadIdx.x >> 5;
    uint32_t acc = 0;
    uint8_t* scratch = smem + 32768 * 3 + 1024;   // 32 KB
    while (!done) {
      if (side & 1) {
        uint32_t r[32];
        tmem_ld_32x32b_x32(tm + ((uint32_t)((warp & 3) * 32) << 16) + 256u + (uint32_t)((warp >> 2) & 3) * 32u, r);
        tmem_ld_wait();
        acc ^= r[3];
      }
      if (side & 2) {
#pragma unroll
        for (int g = 0; g < 4; ++g)
          st_shared_v4(smem_u32(scratch) + (uint32_t)(((threadIdx.x - 64) * 4 + g) * 16) % 32768u, acc, acc, acc, acc);
      }
    }
    if (acc == 0x1234567u) cycles[1] = acc;
  }
  tc_fence_before_sync();
  __syncthreads();
  if (threadIdx.x < 32) tmem_dealloc<512>(tm);
}

int main() {
  long long* c;
  cudaMalloc(&c, 16);
  const int smem = 32768 * 4 + 2048;
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const int reps = 2000;
  const char* names[] = {"S-like M128 N128 KK", "PV-like M128 N128 K,MN", "C-like M128 N16 MN,K", "M128 N256 KK", "M128 N64 KK", "?", "alternate S / C groups", "S-like + commit per group"};
  for (int uniform = 0; uniform < 2; ++uniform)
  for (int side = 0; side < 4; side += 3)
  for (int mode = 0; mode < 8; ++mode) {
    if (mode == 5 || (side && mode > 1 && mode < 6)) continue;
    probe<<<1, side ? 576 : 128, smem>>>(mode, reps, c, side, uniform);
    cudaError_t e = cudaDeviceSynchronize();
    long long h = 0;
    cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
    printf("uniform %d side %d %-24s: %s  %.1f clk per MMA (K = 16), %.1f clk per 8-MMA tile\n", uniform, side, names[mode], cudaGetErrorString(e),
           (double)h / (reps * 8), (double)h / reps);
  }
  return 0;
}
