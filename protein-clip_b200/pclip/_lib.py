"""ctypes binding of libpclip.so (C ABI declared in include/pclip.h).

There is no CPU fallback and no alternative backend: if the shared library is missing the import of any op fails
with a clear message (build it with `make` at the repo root or `python -c "import __graft_entry__ as g; g.build()"`).
"""
from __future__ import annotations

import ctypes as C
import struct
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libpclip.so")

OK = 0
F32, BF16 = 0, 1
COMPUTE_F32, COMPUTE_BF16 = 0, 1
SCORE_DOT, SCORE_EXP = 0, 1


class PclipError(RuntimeError):
    pass


class Chunk(C.Structure):
    _fields_ = [("ptr", C.c_void_p), ("rows", C.c_int64), ("row_stride", C.c_int64), ("col_stride", C.c_int64)]


class Linear(C.Structure):   # pclip_linear_t
    _fields_ = [("weight", C.c_void_p), ("bias", C.c_void_p), ("ln_weight", C.c_void_p), ("ln_bias", C.c_void_p)]


class FwdSide(C.Structure):
    _fields_ = [("a", C.c_void_p), ("b", C.c_void_p), ("Ma", C.c_int64), ("Nb", C.c_int64),
                ("diag_offset", C.c_int64), ("rowsum", C.c_void_p), ("diag", C.c_void_p)]


class BwdSide(C.Structure):
    _fields_ = [("a", C.c_void_p), ("b", C.c_void_p), ("bT", C.c_void_p), ("ldbT", C.c_int64), ("Ma", C.c_int64),
                ("Nb", C.c_int64), ("diag_offset", C.c_int64), ("rinv", C.c_void_p), ("cinv", C.c_void_p),
                ("da", C.c_void_p)]


_P, _I64, _I32, _F, _SZ = C.c_void_p, C.c_int64, C.c_int32, C.c_float, C.c_size_t

# name -> (restype, argtypes); mirrors include/pclip.h one to one
SIGNATURES = {
    "pclip_version": (C.c_int, []),
    "pclip_last_error": (C.c_char_p, []),
    "pclip_launch_count": (_I64, []),
    "pclip_launch_count_reset": (None, []),
    "pclip_normalize_scale_fwd": (C.c_int, [_P, _I64, _I32, _P, _F, _P, _P, _P]),
    "pclip_rowsum_workspace_bytes": (_SZ, [_I64]),
    "pclip_normalize_scale_bwd": (C.c_int, [_P, _P, _P, _I64, _I32, _P, _P, _P, _P, _SZ, _P]),
    "pclip_pack_pair": (C.c_int, [C.POINTER(Chunk), _I32, C.POINTER(Chunk), _I32, _I32, _I64, _I32, _I32, _P, _P, _P]),
    "pclip_pack": (C.c_int, [C.POINTER(Chunk), _I32, _I32, _I64, _I32, _I32, _P, _P, _I64, _P]),
    "pclip_transpose_bf16": (C.c_int, [_P, _I64, _I32, _P, _I64, _P]),
    "pclip_clip_rowsumexp_workspace_bytes": (_SZ, [_I64, _I64, _I32, _I32]),
    "pclip_clip_rowsumexp": (C.c_int, [C.POINTER(FwdSide), _I32, _I32, _I32, _P, _SZ, _P]),
    "pclip_clip_sumexp_workspace_bytes": (_SZ, [_I64, _I64, _I32, _I32]),
    "pclip_clip_sumexp": (C.c_int, [_P, _P, _I64, _I64, _I64, _I32, _I32, _P, _P, _P, _P, _SZ, _P]),
    "pclip_clip_finalize": (C.c_int, [_P, _P, _P, _I64, _F, _P, _P, _P, _P, _P, _P]),
    "pclip_clip_finalize_sharded_workspace_bytes": (_SZ, [_I64]),
    "pclip_clip_finalize_sharded": (C.c_int, [_P, _I32, _I64, _I64, _F, _P, _P, _P, _P, _P, _P, _SZ, _P]),
    "pclip_clip_grad_workspace_bytes": (_SZ, [_I64, _I64, _I32, _I32]),
    "pclip_clip_grad": (C.c_int, [C.POINTER(BwdSide), _I32, _I32, _I32, _F, _P, _P, _SZ, _P]),
    "pclip_clip_workspace_bytes": (_SZ, [_I64, _I32, _I32]),
    "pclip_clip_loss": (C.c_int, [C.POINTER(Chunk), _I32, C.POINTER(Chunk), _I32, _I32, _I64, _I32, _I32, _P, _P, _P,
                                  _P, _P, _P, _SZ, _P]),
    "pclip_clip_small_max_batch": (_I32, []),
    "pclip_clip_small_workspace_bytes": (_SZ, [_I64]),
    "pclip_clip_small_fused": (C.c_int, [C.POINTER(Chunk), _I32, C.POINTER(Chunk), _I32, _I32, _I64, _I32, _P, _P,
                                         C.POINTER(Chunk), C.POINTER(Chunk), _P, _SZ, _P]),
    "pclip_scale_by_scalar": (C.c_int, [_P, _P, _I64, _P, _P]),
    "pclip_scores": (C.c_int, [_P, _P, _I64, _I64, _I32, _I32, _P, _P, _P]),
    "pclip_rank_workspace_bytes": (_SZ, [_I64]),
    "pclip_rank_of_positive": (C.c_int, [_P, _P, _P, _I64, _I64, _I32, _P, _P, _SZ, _P]),
    "pclip_topk_bank_bytes": (_SZ, [_I64, _I32]),
    "pclip_topk_prepare_bank": (C.c_int, [_P, _I64, _I32, _P, _SZ, _P]),
    "pclip_topk_prepared": (C.c_int, [_P, _P, _P, _I64, _I64, _I32, _I32, _I64, _P, _P, _P, _P, _SZ, _P]),
    "pclip_topk_workspace_bytes": (_SZ, [_I64, _I64, _I32, _I32, _I32]),
    "pclip_topk": (C.c_int, [_P, _P, _I64, _I64, _I32, _I32, _I32, _I64, _P, _P, _P, _P, _SZ, _P]),
    "pclip_topk_merge": (C.c_int, [_P, _P, _I64, _I32, _I32, _P, _P, _P]),
    "pclip_filip_workspace_bytes": (_SZ, [_I64, _I64, _I64, _I64, _I32, _I32]),
    "pclip_filip_fwd": (C.c_int, [_P, _P, _P, _P, _I64, _I64, _I64, _I64, _I32, _P, _I32, _P, _P, _P, _P, _P, _SZ, _P]),
    "pclip_filip_bwd": (C.c_int, [_P, _P, _P, _P, _I64, _I64, _I64, _I64, _I32, _P, _I32, _P, _P, _P, _P, _P, _P, _P, _P,
                                  _P, _P, _SZ, _P]),
    "pclip_encoder_head_workspace_bytes": (_SZ, [_I64, _I64]),
    "pclip_encoder_head_fwd": (C.c_int, [_P, _P, _I64, _I64, _I32, _I32, C.POINTER(Linear), C.POINTER(Linear), _I32,
                                         C.POINTER(Linear), _I32, _P, _F, _F, _P, _P, _P, _SZ, _P]),
    "pclip_rowwise_infonce_noalloc_max_rows": (_I64, []),
    "pclip_rowwise_infonce_noalloc": (C.c_int, [_P, _I64, _I64, _I64, _F, _P, _P, _P]),
    "pclip_rowwise_infonce": (C.c_int, [_P, _I64, _I64, _I64, _F, _P, _P, _P, _SZ, _P]),
}

_lib = None


def lib():
    """The loaded library (loaded once).  Raises PclipError when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.isfile(LIB_PATH):
            raise PclipError(
                f"{LIB_PATH} not found: the CUDA extension is not built (run `make` in the repo root). "
                "protein-clip_b200 has no CPU fallback.")
        h = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(h, name)  # AttributeError if the library does not export a declared symbol
            fn.restype = res
            fn.argtypes = args
        _lib = h
    return _lib


def check(rc: int):
    if rc != OK:
        raise PclipError(f"libpclip error {rc}: {lib().pclip_last_error().decode(errors='replace')}")


def stream_ptr(device=None) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def ptr(t):
    return None if t is None else t.data_ptr()


def workspace(nbytes: int, device) -> tuple[torch.Tensor, int]:
    """(owner tensor, 1024-byte aligned device pointer) with at least nbytes behind the pointer."""
    buf = torch.empty(int(nbytes) + 1024, dtype=torch.uint8, device=device)
    p = (buf.data_ptr() + 1023) & ~1023
    return buf, p


def dtype_code(dt: torch.dtype) -> int:
    if dt == torch.float32:
        return F32
    if dt == torch.bfloat16:
        return BF16
    raise PclipError(f"unsupported dtype {dt}")


def chunks_of(tensors) -> "C.Array[Chunk]":
    # one packed buffer instead of a ctypes Structure per tensor: this runs once per leaf on the GradCache path
    flat = []
    for t in tensors:
        assert t.dim() == 2
        st = t.stride()
        flat += (t.data_ptr(), t.shape[0], st[0], st[1])
    return (Chunk * len(tensors)).from_buffer_copy(struct.pack(f"{len(flat)}q", *flat))


def require_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise PclipError("protein-clip_b200 ops need CUDA tensors: there is no CPU fallback "
                             f"(got a tensor on {t.device})")
