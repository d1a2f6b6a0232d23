"""torch-facing wrappers of the libpclip.so entry points (device memory, streams and autograd plumbing only).

Every function here takes CUDA tensors and calls the C ABI; none of them computes on the CPU.
"""
from __future__ import annotations

import os

import torch

from . import _lib as L

F32 = L.COMPUTE_F32
BF16 = L.COMPUTE_BF16


def default_compute() -> int:
    v = os.environ.get("PCLIP_COMPUTE", "").lower()
    if v in ("bf16", "1"):
        return BF16
    return F32


def launch_count() -> int:
    return int(L.lib().pclip_launch_count())


def reset_launch_count():
    L.lib().pclip_launch_count_reset()


# ---------------------------------------------------------------------------------------------------------
# (1) encoder tail: F.normalize(e) * exp(tau / 2)      ref: modules/models.py:50-51
# ---------------------------------------------------------------------------------------------------------
class _NormalizeScale(torch.autograd.Function):
    @staticmethod
    def forward(ctx, e, tau, eps):
        L.require_cuda(e)
        shape = e.shape
        e2 = e.reshape(-1, shape[-1]).contiguous().float()
        rows, D = e2.shape
        out = torch.empty_like(e2)
        inv = torch.empty(rows, dtype=torch.float32, device=e.device)
        tau_d = None if tau is None else tau.detach().reshape(1).float().to(e.device)
        L.check(L.lib().pclip_normalize_scale_fwd(L.ptr(e2), rows, D, L.ptr(tau_d), float(eps), L.ptr(out), L.ptr(inv),
                                                  L.stream_ptr(e.device)))
        ctx.save_for_backward(out, inv, tau_d if tau_d is not None else torch.empty(0, device=e.device))
        ctx.has_tau = tau is not None
        ctx.shape = shape
        ctx.tau_shape = None if tau is None else tau.shape
        return out.reshape(shape)

    @staticmethod
    def backward(ctx, g):
        out, inv, tau_d = ctx.saved_tensors
        rows, D = out.shape
        g2 = g.reshape(rows, D).contiguous().float()
        de = torch.empty_like(out)
        dtau = torch.zeros(1, dtype=torch.float32, device=out.device) if ctx.has_tau else None
        L.check(L.lib().pclip_normalize_scale_bwd(L.ptr(g2), L.ptr(out), L.ptr(inv), rows, D,
                                                  L.ptr(tau_d) if ctx.has_tau else None, L.ptr(de), L.ptr(dtau),
                                                  L.stream_ptr(out.device)))
        return de.reshape(ctx.shape), (dtau.reshape(ctx.tau_shape) if ctx.has_tau else None), None


def normalize_scale(e: torch.Tensor, tau: torch.Tensor | None, eps: float = 1e-12) -> torch.Tensor:
    """x / max(||x||, eps) * exp(tau / 2) over the last dim (tau=None, eps=0: FILIP's h / h.norm())."""
    return _NormalizeScale.apply(e, tau, eps)


# ---------------------------------------------------------------------------------------------------------
# (2) CLIP InfoNCE                                        ref: modules/training_utils.py:74-89, :247-265
# ---------------------------------------------------------------------------------------------------------
def _as_rows(ts):
    out = []
    for t in ts:
        if t.dtype not in (torch.float32, torch.bfloat16):
            t = t.float()
        out.append(t)
    dt = out[0].dtype
    return [t if t.dtype == dt else t.to(dt) for t in out]


def clip_loss_raw(xs, ys, compute: int, want_grad: bool, want_lse: bool = False):
    """xs / ys: lists of [rows_i, D] CUDA tensors (any strides).  Returns (loss[1], dx, dy, row_lse, col_lse) with
    dx / dy packed [B, D] fp32 for grad_output == 1 (None unless want_grad)."""
    L.require_cuda(*xs, *ys)
    xs, ys = _as_rows(xs), _as_rows(ys)
    if ys[0].dtype != xs[0].dtype:
        ys = [t.to(xs[0].dtype) for t in ys]
    dev = xs[0].device
    B = sum(t.shape[0] for t in xs)
    D = xs[0].shape[1]
    if sum(t.shape[0] for t in ys) != B or any(t.shape[1] != D for t in xs + ys):
        raise L.PclipError("clip_loss: x and y must hold the same number of [*, D] rows")
    lib = L.lib()
    nbytes = lib.pclip_clip_workspace_bytes(B, D, compute)
    owner, ws = L.workspace(nbytes, dev)
    loss = torch.empty(1, dtype=torch.float32, device=dev)
    dx = torch.empty(B, D, dtype=torch.float32, device=dev) if want_grad else None
    dy = torch.empty(B, D, dtype=torch.float32, device=dev) if want_grad else None
    rl = torch.empty(B, dtype=torch.float32, device=dev) if want_lse else None
    cl = torch.empty(B, dtype=torch.float32, device=dev) if want_lse else None
    with torch.cuda.device(dev):
        L.check(lib.pclip_clip_loss(L.chunks_of(xs), len(xs), L.chunks_of(ys), len(ys), L.dtype_code(xs[0].dtype), B, D,
                                    compute, L.ptr(loss), L.ptr(rl), L.ptr(cl), L.ptr(dx), L.ptr(dy), ws, nbytes,
                                    L.stream_ptr(dev)))
    owner.record_stream(torch.cuda.current_stream(dev))
    return loss, dx, dy, rl, cl


class _ClipInfoNCE(torch.autograd.Function):
    """loss = ClipInfoNCE(nx, compute, *x_chunks, *y_chunks).  Gradients are produced together with the loss (one
    fused pass over the score tiles) and scaled by grad_output on the device in backward, so GradScaler's 65536
    (ref: modules/training_utils.py:139) never needs a host sync."""

    @staticmethod
    def forward(ctx, nx, compute, *tensors):
        xs, ys = list(tensors[:nx]), list(tensors[nx:])
        need = any(t.requires_grad for t in tensors)
        loss, dx, dy, _, _ = clip_loss_raw([t.detach() for t in xs], [t.detach() for t in ys], compute, need)
        ctx.nx = nx
        ctx.rows = [t.shape[0] for t in tensors]
        ctx.dtypes = [t.dtype for t in tensors]
        if need:
            ctx.save_for_backward(dx, dy)
        return loss.reshape(())

    @staticmethod
    def backward(ctx, g):
        dx, dy = ctx.saved_tensors
        g = g.reshape(1).float().contiguous()
        lib = L.lib()
        outs = []
        for src in (dx, dy):
            o = torch.empty_like(src)
            L.check(lib.pclip_scale_by_scalar(L.ptr(src), L.ptr(g), src.numel(), L.ptr(o), L.stream_ptr(src.device)))
            outs.append(o)
        grads = []
        r = 0
        for i, n in enumerate(ctx.rows):
            if i == ctx.nx:
                r = 0
            src = outs[0] if i < ctx.nx else outs[1]
            piece = src[r:r + n]
            grads.append(piece if ctx.dtypes[i] == torch.float32 else piece.to(ctx.dtypes[i]))
            r += n
        return (None, None, *grads)


def clip_loss(x, y, compute: int | None = None) -> torch.Tensor:
    """Symmetric InfoNCE of S = x @ y.T.  x / y: [B, D] tensors (any strides) or lists of [rows_i, D] tensors."""
    xs = list(x) if isinstance(x, (list, tuple)) else [x]
    ys = list(y) if isinstance(y, (list, tuple)) else [y]
    if compute is None:
        compute = default_compute()
    return _ClipInfoNCE.apply(len(xs), compute, *xs, *ys)
