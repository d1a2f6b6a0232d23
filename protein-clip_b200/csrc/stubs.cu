// Entry points declared in include/pclip.h whose kernels are not written yet: they fail loudly.
#include "common.cuh"
using namespace pclip;
extern "C" {
size_t pclip_filip_workspace_bytes(int64_t, int64_t, int64_t, int64_t, int32_t, int32_t) { return 0; }
int pclip_filip_fwd(const float*, const float*, const int32_t*, const int32_t*, int64_t, int64_t, int64_t, int64_t,
                    int32_t, const float*, int32_t, float*, float*, int16_t*, int16_t*, void*, size_t, pclip_stream_t) {
  return fail(PCLIP_ERR_UNSUPPORTED, "pclip_filip_fwd: not implemented");
}
int pclip_filip_bwd(const float*, const float*, const int32_t*, const int32_t*, int64_t, int64_t, int64_t, int64_t,
                    int32_t, const float*, const float*, const float*, const int16_t*, const int16_t*, const float*,
                    const float*, float*, float*, float*, void*, size_t, pclip_stream_t) {
  return fail(PCLIP_ERR_UNSUPPORTED, "pclip_filip_bwd: not implemented");
}
}
