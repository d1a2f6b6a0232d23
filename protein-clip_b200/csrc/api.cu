// libpclip.so: error plumbing, launch bookkeeping and the composite CLIP entry points (include/pclip.h).
#include <stdlib.h>
#include <string.h>

#include "common.cuh"

namespace pclip {

static thread_local char g_err[512] = "";
static thread_local int64_t g_launches = 0;

int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

void count_launch(int n) { g_launches += n; }

int device_sm_count() {
  static thread_local int cached_dev = -1;
  static thread_local int cached_sms = 0;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 0;
  if (dev != cached_dev) {
    cudaDeviceProp p;
    if (cudaGetDeviceProperties(&p, dev) != cudaSuccess) return 0;
    cached_dev = dev;
    cached_sms = p.multiProcessorCount;
  }
  return cached_sms;
}

int require_device() {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0) {
    cudaGetLastError();
    return fail(PCLIP_ERR_CUDA, "no CUDA device: libpclip has no CPU fallback (%s)",
                e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
  }
  return PCLIP_OK;
}

}  // namespace pclip

using namespace pclip;

extern "C" {

int pclip_version(void) { return 100; }
const char* pclip_last_error(void) { return g_err; }
int64_t pclip_launch_count(void) { return g_launches; }
void pclip_launch_count_reset(void) { g_launches = 0; }

size_t pclip_clip_rowsumexp_workspace_bytes(int64_t Ma, int64_t Nb, int32_t D, int32_t compute) {
  (void)D;
  return compute == PCLIP_COMPUTE_BF16 ? tc_rowsumexp_workspace(Ma, Nb) : 0;
}

size_t pclip_clip_grad_workspace_bytes(int64_t Ma, int64_t Nb, int32_t D, int32_t compute) {
  (void)D;
  if (compute != PCLIP_COMPUTE_BF16) return 0;
  return tc2_grad_workspace(Ma, Nb);
}

static size_t ld_transposed(int64_t rows) { return align_up((size_t)rows, 64); }

size_t pclip_clip_workspace_bytes(int64_t B, int32_t D, int32_t compute) {
  // packed X, Y, 8 per-row vectors, kernel scratch, alignment slack
  const bool tc = compute == PCLIP_COMPUTE_BF16;
  const size_t esz = tc ? 2 : 4;
  size_t n = 2 * align_up((size_t)B * D * esz, 1024) + 8 * align_up((size_t)B * sizeof(float), 1024);
  if (tc) {
    const size_t a = std::max(tc_rowsumexp_workspace(B, B), tc2_sumexp_workspace(B, B));
    const size_t b = tc2_grad_workspace(B, B);
    n += align_up(a > b ? a : b, 1024);
  }
  return n + 16 * 1024;
}

int pclip_clip_rowsumexp(const pclip_clip_fwd_side_t* sides, int32_t n_sides, int32_t D, int32_t compute,
                         void* workspace, size_t workspace_bytes, pclip_stream_t stream) {
  PCLIP_TRY(require_device());
  PCLIP_REQUIRE(sides && n_sides >= 1 && n_sides <= 2, PCLIP_ERR_INVALID, "rowsumexp: n_sides must be 1 or 2");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (compute == PCLIP_COMPUTE_BF16) return tc_rowsumexp(sides, n_sides, D, workspace, workspace_bytes, st);
  if (compute == PCLIP_COMPUTE_F32) return simt_rowsumexp(sides, n_sides, D, st);
  return fail(PCLIP_ERR_UNSUPPORTED, "rowsumexp: unknown compute mode %d", compute);
}

size_t pclip_clip_sumexp_workspace_bytes(int64_t Ma, int64_t Nb, int32_t D, int32_t compute) {
  (void)D;
  return compute == PCLIP_COMPUTE_BF16 ? tc2_sumexp_workspace(Ma, Nb) : 0;
}

int pclip_clip_sumexp(const void* a, const void* b, int64_t Ma, int64_t Nb, int64_t diag_offset, int32_t D,
                      int32_t compute, float* rowsum, float* colsum, float* diag, void* workspace,
                      size_t workspace_bytes, pclip_stream_t stream) {
  PCLIP_TRY(require_device());
  PCLIP_REQUIRE(a && b && rowsum && colsum && diag && Ma > 0 && Nb > 0, PCLIP_ERR_INVALID, "sumexp: null argument");
  PCLIP_REQUIRE(diag_offset >= 0 && diag_offset + Ma <= Nb, PCLIP_ERR_INVALID, "sumexp: positives fall outside b");
  PCLIP_REQUIRE(compute == PCLIP_COMPUTE_BF16, PCLIP_ERR_UNSUPPORTED,
                "sumexp: single-pass forward exists for bf16 compute only (use pclip_clip_rowsumexp)");
  return tc2_sumexp(a, b, Ma, Nb, diag_offset, D, rowsum, colsum, diag, workspace, workspace_bytes,
                    static_cast<cudaStream_t>(stream));
}

int pclip_clip_grad(const pclip_clip_bwd_side_t* sides, int32_t n_sides, int32_t D, int32_t compute, float scale,
                    const float* gscale, void* workspace, size_t workspace_bytes, pclip_stream_t stream) {
  PCLIP_TRY(require_device());
  PCLIP_REQUIRE(sides && n_sides >= 1 && n_sides <= 2, PCLIP_ERR_INVALID, "grad: n_sides must be 1 or 2");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (compute == PCLIP_COMPUTE_BF16)
    return tc2_grad(sides, n_sides, D, scale, gscale, workspace, workspace_bytes, st);
  if (compute == PCLIP_COMPUTE_F32) return simt_grad(sides, n_sides, D, scale, gscale, st);
  return fail(PCLIP_ERR_UNSUPPORTED, "grad: unknown compute mode %d", compute);
}

int pclip_clip_loss(const pclip_chunk_t* x_chunks, int32_t nx, const pclip_chunk_t* y_chunks, int32_t ny,
                    int32_t in_dtype, int64_t B, int32_t D, int32_t compute, float* loss, float* row_lse,
                    float* col_lse, float* dx, float* dy, void* workspace, size_t workspace_bytes,
                    pclip_stream_t stream) {
  PCLIP_TRY(require_device());
  PCLIP_REQUIRE(x_chunks && y_chunks && nx > 0 && ny > 0 && loss, PCLIP_ERR_INVALID, "clip_loss: null argument");
  PCLIP_REQUIRE(B > 0 && D > 0, PCLIP_ERR_INVALID, "clip_loss: B and D must be positive");
  PCLIP_REQUIRE((dx == nullptr) == (dy == nullptr), PCLIP_ERR_INVALID, "clip_loss: dx and dy go together");
  PCLIP_REQUIRE(compute == PCLIP_COMPUTE_F32 || compute == PCLIP_COMPUTE_BF16, PCLIP_ERR_UNSUPPORTED,
                "clip_loss: unknown compute mode %d", compute);
  PCLIP_REQUIRE(workspace && (reinterpret_cast<uintptr_t>(workspace) & 1023) == 0, PCLIP_ERR_WORKSPACE,
                "clip_loss: workspace must be 1024-byte aligned");
  PCLIP_REQUIRE(workspace_bytes >= pclip_clip_workspace_bytes(B, D, compute), PCLIP_ERR_WORKSPACE,
                "clip_loss: workspace too small (%zu < %zu)", workspace_bytes,
                pclip_clip_workspace_bytes(B, D, compute));
  const bool tc = compute == PCLIP_COMPUTE_BF16;
  const int cdt = tc ? PCLIP_BF16 : PCLIP_F32;
  const size_t esz = tc ? 2 : 4;
  const int64_t ldT = (int64_t)ld_transposed(B);
  const bool want_grad = dx != nullptr;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  Workspace ws(workspace, workspace_bytes);
  char* X = ws.take<char>((size_t)B * D * esz);
  char* Y = ws.take<char>((size_t)B * D * esz);
  char* XT = nullptr;
  char* YT = nullptr;
  // one record [C | R | S_ii], the layout pclip_clip_finalize_sharded reads (P = 1)
  float* stats = ws.take<float>(3 * (size_t)B);
  float* colsum = stats;
  float* rowsum = stats + B;
  float* diag = stats + 2 * B;
  float* diag2 = ws.take<float>(B);
  float* rinv = ws.take<float>(B);
  float* cinv = ws.take<float>(B);
  float* rl = row_lse ? row_lse : ws.take<float>(B);
  float* cl = col_lse ? col_lse : ws.take<float>(B);
  ws.off = align_up(ws.off, 1024);
  void* scratch = ws.base + ws.off;
  const size_t scratch_bytes = ws.size > ws.off ? ws.size - ws.off : 0;
  PCLIP_REQUIRE(ws.ok(), PCLIP_ERR_WORKSPACE, "clip_loss: workspace overflow");

  PCLIP_TRY(pack_operands(x_chunks, nx, y_chunks, ny, in_dtype, B, D, cdt, X, XT, Y, YT, ldT, st));
  // below 2048 columns the two-sided forward keeps the column sums fp32-exact (the single pass sums bf16-rounded
  // exps: the rounding averages out over thousands of rows, not over 16)
  if (tc && B >= 2048) {
    PCLIP_TRY(pclip_clip_sumexp(X, Y, B, B, 0, D, compute, rowsum, colsum, diag, scratch, scratch_bytes, stream));
  } else {
    pclip_clip_fwd_side_t f[2] = {{X, Y, B, B, 0, rowsum, diag}, {Y, X, B, B, 0, colsum, diag2}};
    PCLIP_TRY(pclip_clip_rowsumexp(f, 2, D, compute, scratch, scratch_bytes, stream));
  }
  const float scale = 0.5f / (float)B;
  if (B >= 4096 && scratch_bytes >= pclip_clip_finalize_sharded_workspace_bytes(B)) {
    PCLIP_TRY(pclip_clip_finalize_sharded(stats, 1, B, B, scale, rl, cl, rinv, cinv, loss, scratch, scratch_bytes, stream));
  } else {
    PCLIP_TRY(pclip_clip_finalize(rowsum, colsum, diag, B, scale, rl, cl, rinv, cinv, loss, stream));
  }
  if (want_grad) {
    pclip_clip_bwd_side_t g[2] = {{X, Y, YT, ldT, B, B, 0, rinv, cinv, dx}, {Y, X, XT, ldT, B, B, 0, cinv, rinv, dy}};
    PCLIP_TRY(pclip_clip_grad(g, 2, D, compute, scale, nullptr, scratch, scratch_bytes, stream));
  }
  return PCLIP_OK;
}

}  // extern "C"
