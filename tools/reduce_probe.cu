// Probe for the single-S CLIP backward: how fast can 148 SMs add fp32 tile partials into an
// L2-resident [65536,128] fp32 gradient?  A single-S gradient kernel forms P once per (row panel, column tile) and must
// add a 128x128 fp32 dY partial (64 KB) per tile into global memory: 8.6 GB of reduce traffic per step with 256-row
// panels, 17.2 GB with 128-row panels.  The current two-sided kernel (tc3_grad_kernel, 3.52 ms) pays one extra S-MMA
// per tile instead (+0.85 ms of tensor time).  This binary measures the reduce path alone, three ways:
//   bulk : cp.reduce.async.bulk.global.shared::cta.add.f32, one 64 KB op per tile from shared memory (TMA engine)
//   red  : red.global.add.v4.f32 from registers, fully coalesced (16 B per thread)
//   st   : plain st.global.v4.f32 of the same bytes (what the L2 takes with no read-modify-write)
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o reduce_probe tools/reduce_probe.cu ; ./reduce_probe
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); return 1; } } while (0)

constexpr int kTileBytes = 64 * 1024;      // 128 rows x 128 fp32
constexpr int kTiles = 512;                // 65536 / 128 row tiles of the gradient
constexpr int kThreads = 256;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <int MODE>
__global__ void __launch_bounds__(kThreads, 1) probe_kernel(float* __restrict__ dst, int tiles_per_cta, int stride) {
    extern __shared__ __align__(128) unsigned char smem[];
    float4* tile = reinterpret_cast<float4*>(smem);
    for (int i = threadIdx.x; i < 2 * kTileBytes / 16; i += kThreads) tile[i] = make_float4(1.f, 1.f, 1.f, 1.f);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    int t = (blockIdx.x * stride) % kTiles;
    for (int it = 0; it < tiles_per_cta; ++it) {
        char* g = reinterpret_cast<char*>(dst) + (size_t)t * kTileBytes;
        if (MODE == 0) {
            if (threadIdx.x == 0) {
                // double buffered: at most two 64 KB reduce ops in flight per SM
                asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], %2;"
                             :: "l"(g), "r"(smem_u32(smem + (it & 1) * kTileBytes)), "r"(kTileBytes) : "memory");
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
            }
        } else {
            float4 v = tile[threadIdx.x];
            float4* gp = reinterpret_cast<float4*>(g);
#pragma unroll 4
            for (int i = threadIdx.x; i < kTileBytes / 16; i += kThreads) {
                if (MODE == 1)
                    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};"
                                 :: "l"(gp + i), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
                else
                    asm volatile("st.global.v4.f32 [%0], {%1, %2, %3, %4};"
                                 :: "l"(gp + i), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
            }
        }
        t += 1;
        if (t >= kTiles) t -= kTiles;
    }
    if (MODE == 0 && threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

template <int MODE>
static int run(const char* name, float* dst, int ctas, int tiles_per_cta, int stride) {
    CK(cudaFuncSetAttribute(probe_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * kTileBytes));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep) {
        CK(cudaMemsetAsync(dst, 0, (size_t)kTiles * kTileBytes));
        CK(cudaEventRecord(e0));
        probe_kernel<MODE><<<ctas, kThreads, 2 * kTileBytes>>>(dst, tiles_per_cta, stride);
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        CK(cudaGetLastError());
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        if (rep && ms < best) best = ms;
    }
    double bytes = (double)ctas * tiles_per_cta * kTileBytes;
    float h[4];
    CK(cudaMemcpy(h, dst, sizeof(h), cudaMemcpyDeviceToHost));
    printf("{\"mode\": \"%s\", \"ctas\": %d, \"tiles_per_cta\": %d, \"stride\": %d, \"gbytes\": %.3f, \"ms\": %.4f, "
           "\"tb_per_s\": %.3f, \"dst0\": %.1f}\n", name, ctas, tiles_per_cta, stride, bytes / 1e9, best,
           bytes / best / 1e9, h[0]);
    return 0;
}

int main() {
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    float* dst;
    CK(cudaMalloc(&dst, (size_t)kTiles * kTileBytes));
    // 8.6 GB = 131072 tile partials (512 column tiles x 256 row panels of 256 rows): 886 per SM
    int per = (512 * 256 + sms - 1) / sms;
    // stride 3: neighbouring SMs work on different tiles, every tile is hit by all SMs over time (as the sweep would)
    if (run<0>("bulk", dst, sms, per, 3)) return 1;
    if (run<1>("red", dst, sms, per, 3)) return 1;
    if (run<2>("st", dst, sms, per, 3)) return 1;
    // all SMs on the same tile at the same time (a column-major sweep without staggering): per-address serialisation
    if (run<0>("bulk_same_tile", dst, sms, per / 4, 0)) return 1;
    if (run<1>("red_same_tile", dst, sms, per / 4, 0)) return 1;
    cudaFree(dst);
    return 0;
}
