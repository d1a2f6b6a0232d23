// TMA tensor-map construction.  cuTensorMapEncodeTiled is fetched through the runtime's driver entry point so
// that libpclip.so has no link-time dependency on libcuda.so (absent in the build container).
#include <cudaTypedefs.h>

#include "common.cuh"

namespace pclip {

static PFN_cuTensorMapEncodeTiled_v12000 encode_fn() {
  static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(p);
  }
  return fn;
}

int make_tmap_bf16_2d(CUtensorMap* map, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld_elems,
                      uint32_t box_rows, uint32_t box_cols) {
  auto fn = encode_fn();
  PCLIP_REQUIRE(fn, PCLIP_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
  PCLIP_REQUIRE((reinterpret_cast<uintptr_t>(ptr) & 15) == 0 && (ld_elems * 2) % 16 == 0, PCLIP_ERR_INVALID,
                "tensor map: base / row pitch must be 16-byte aligned");
  cuuint64_t gdim[2] = {cols, rows};                  // innermost first
  cuuint64_t gstride[1] = {ld_elems * 2};             // bytes, dims 1..
  cuuint32_t box[2] = {box_cols, box_rows};           // 64 bf16 = 128 B = one swizzle atom row
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), gdim, gstride, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  PCLIP_REQUIRE(r == CUDA_SUCCESS, PCLIP_ERR_CUDA, "cuTensorMapEncodeTiled(2d) failed: CUresult %d", (int)r);
  return PCLIP_OK;
}

int make_tmap_bf16_3d(CUtensorMap* map, const void* ptr, uint64_t batch, uint64_t rows, uint64_t cols,
                      uint32_t box_rows, uint32_t box_cols) {
  auto fn = encode_fn();
  PCLIP_REQUIRE(fn, PCLIP_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
  cuuint64_t gdim[3] = {cols, rows, batch};
  cuuint64_t gstride[2] = {cols * 2, rows * cols * 2};
  cuuint32_t box[3] = {box_cols, box_rows, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(ptr), gdim, gstride, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  PCLIP_REQUIRE(r == CUDA_SUCCESS, PCLIP_ERR_CUDA, "cuTensorMapEncodeTiled(3d) failed: CUresult %d", (int)r);
  return PCLIP_OK;
}

}  // namespace pclip
