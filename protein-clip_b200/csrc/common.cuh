// Host-side plumbing shared by every translation unit of libpclip.so.
#pragma once

#include <cuda.h>
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>

#include <algorithm>

#include "../../include/pclip.h"

namespace pclip {

// thread-local error text + launch counter (defined in api.cu)
int fail(int code, const char* fmt, ...);
void count_launch(int n = 1);

#define PCLIP_CHECK_CUDA(expr)                                                                        \
  do {                                                                                                \
    cudaError_t _e = (expr);                                                                          \
    if (_e != cudaSuccess)                                                                            \
      return ::pclip::fail(PCLIP_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e),    \
                           __FILE__, __LINE__);                                                       \
  } while (0)

#define PCLIP_CHECK_LAUNCH()                                                                          \
  do {                                                                                                \
    cudaError_t _e = cudaGetLastError();                                                              \
    if (_e != cudaSuccess)                                                                            \
      return ::pclip::fail(PCLIP_ERR_CUDA, "kernel launch failed: %s (%s:%d)", cudaGetErrorString(_e),\
                           __FILE__, __LINE__);                                                       \
    ::pclip::count_launch();                                                                          \
  } while (0)

#define PCLIP_REQUIRE(cond, code, ...)                                                                \
  do {                                                                                                \
    if (!(cond)) return ::pclip::fail(code, __VA_ARGS__);                                             \
  } while (0)

#define PCLIP_TRY(expr)                                                                               \
  do {                                                                                                \
    int _r = (expr);                                                                                  \
    if (_r != PCLIP_OK) return _r;                                                                    \
  } while (0)

__host__ __device__ inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }
inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// bump allocator over the caller's workspace
struct Workspace {
  char* base;
  size_t size;
  size_t off = 0;
  Workspace(void* p, size_t n) : base(static_cast<char*>(p)), size(n) {}
  template <class T>
  T* take(size_t count) {
    off = align_up(off, 1024);
    T* r = reinterpret_cast<T*>(base + off);
    off += count * sizeof(T);
    return r;
  }
  bool ok() const { return off <= size; }
};

// 2-D bf16 row-major [rows, cols] tensor map with a [box_rows, 64] box and 128-byte swizzle (tma.cu)
int make_tmap_bf16_2d(CUtensorMap* map, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld_elems,
                      uint32_t box_rows, uint32_t box_cols);
// 3-D bf16 [d0=batch, d1=tokens, d2=cols] (cols innermost), box [1, box_rows, 64]
int make_tmap_bf16_3d(CUtensorMap* map, const void* ptr, uint64_t batch, uint64_t rows, uint64_t cols,
                      uint32_t box_rows, uint32_t box_cols);

int device_sm_count();
// PCLIP_OK, or PCLIP_ERR_CUDA when no device is visible (there is no CPU fallback)
int require_device();

// elementwise.cu: packs one or two operand chunk lists (same shape) with one launch
int pack_operands(const pclip_chunk_t* x, int nx, const pclip_chunk_t* y, int ny, int in_dtype, int64_t rows, int D,
                  int out_dtype, void* outx, void* outxT, void* outy, void* outyT, int64_t ldT, cudaStream_t st);

// ---- kernel-family entry points (one per .cu) ----------------------------------------------
// FP32 CUDA-core path (clip_simt.cu)
int simt_rowsumexp(const pclip_clip_fwd_side_t* sides, int n_sides, int D, cudaStream_t st);
int simt_grad(const pclip_clip_bwd_side_t* sides, int n_sides, int D, float scale, const float* gscale,
              cudaStream_t st);
// BF16 tcgen05 path (clip_tc.cu)
size_t tc_rowsumexp_workspace(int64_t Ma, int64_t Nb);
int tc_rowsumexp(const pclip_clip_fwd_side_t* sides, int n_sides, int D, void* ws, size_t ws_bytes,
                 cudaStream_t st);
// one side of the reduction of the gradient kernel's column-split partials (clip_tc.cu: grad_reduce_kernel)
struct GradReduceSide {
  const float* partial;       // [nsplit][Ma][128]
  const void* b;              // packed bf16 [Nb, 128] (the -2 b_pos term)
  float* da;                  // [Ma, 128]
  int64_t diag_offset;
};

// single-pass forward and gradient kernels (clip_tc2.cu)
size_t tc2_grad_workspace(int64_t Ma, int64_t Nb);
int tc2_grad(const pclip_clip_bwd_side_t* sides, int n_sides, int D, float scale, const float* gscale, void* ws,
             size_t ws_bytes, cudaStream_t st);
size_t tc2_sumexp_workspace(int64_t Ma, int64_t Nb);
int tc2_sumexp(const void* a, const void* b, int64_t Ma, int64_t Nb, int64_t diag_offset, int D, float* rowsum,
               float* colsum, float* diag, void* ws, size_t ws_bytes, cudaStream_t st);
// BF16 tcgen05 candidate selection + fp32 re-scoring (topk_tc.cu)
size_t tc_topk_workspace(int64_t Q, int64_t N, int k);
int tc_topk(const float* q, const float* bank, int64_t Q, int64_t N, int D, int k, int64_t index_base, int64_t* idx,
            float* score, int32_t* n_unsafe, void* ws, size_t ws_bytes, cudaStream_t st,
            const void* prepared = nullptr);
size_t tc_topk_bank_bytes(int64_t N);
int tc_topk_prepare_bank(const float* bank, int64_t N, int D, void* prepared, size_t prepared_bytes, cudaStream_t st);
// FILIP (filip.cu: fp32 path + tails; filip_tc.cu: bf16 tcgen05 path)
int filip_finalize(const float* part, int np, const int32_t* lenX, int Bx, int By, int Tx, const float* tau, float* out,
                   int64_t so0, int64_t so1, cudaStream_t st);
size_t tc_filip_workspace(int64_t Ba, int64_t Bb, int64_t TA, int64_t TB);
int tc_filip_fwd(const float* hA, const float* hB, const int32_t* lenA, const int32_t* lenB, int64_t Ba, int64_t Bb,
                 int64_t TA, int64_t TB, int D, const float* tau, float* simA, float* simB, int16_t* argA,
                 int16_t* argB, void* ws, size_t ws_bytes, cudaStream_t st);
int tc_filip_bwd(const float* hA, const float* hB, const int32_t* lenA, const int32_t* lenB, int64_t Ba, int64_t Bb,
                 int64_t TA, int64_t TB, int D, const float* tau, const int16_t* argA, const int16_t* argB,
                 const float* gA, const float* gB, float* dhA, float* dhB, void* ws, size_t ws_bytes, cudaStream_t st);
// exact fp32 top-k on CUDA cores (scores.cu)
// qlist / qcount (device, optional): "list mode" -- only queries qlist[0 .. *qcount) are recomputed, in place
size_t simt_topk_workspace(int64_t Q, int64_t N, int k, bool list_mode);
int simt_topk(const float* q, const float* bank, int64_t Q, int64_t N, int D, int k, int64_t index_base, int64_t* idx,
              float* score, void* ws, size_t ws_bytes, cudaStream_t st, const int32_t* qlist = nullptr,
              const int32_t* qcount = nullptr);

}  // namespace pclip
