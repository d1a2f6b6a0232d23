"""torch-facing wrappers of the libpclip.so entry points (device memory, streams and autograd plumbing only).

Every function here takes CUDA tensors and calls the C ABI; none of them computes on the CPU.
"""
from __future__ import annotations

import os

import torch

from . import _lib as L

F32 = L.COMPUTE_F32
BF16 = L.COMPUTE_BF16
SINGLE_PASS_MIN_COLS = 2048   # same threshold as pclip_clip_loss (csrc/api.cu)


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
        with torch.cuda.device(e.device):
            L.check(L.lib().pclip_normalize_scale_fwd(L.ptr(e2), rows, D, L.ptr(tau_d), float(eps), L.ptr(out),
                                                      L.ptr(inv), L.stream_ptr(e.device)))
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
        with torch.cuda.device(out.device):
            nbytes = int(L.lib().pclip_rowsum_workspace_bytes(rows)) if ctx.has_tau else 0
            owner, ws = L.workspace(nbytes, out.device) if ctx.has_tau else (None, None)
            L.check(L.lib().pclip_normalize_scale_bwd(L.ptr(g2), L.ptr(out), L.ptr(inv), rows, D,
                                                      L.ptr(tau_d) if ctx.has_tau else None, L.ptr(de), L.ptr(dtau),
                                                      ws, nbytes, L.stream_ptr(out.device)))
        if owner is not None:
            owner.record_stream(torch.cuda.current_stream(out.device))
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


MAX_CHUNKS = 32   # pclip_pack takes at most 32 row chunks per modality (csrc/elementwise.cu kMaxChunks)


def _limit_chunks(ts):
    """More than 32 leaves (train_gc with accumulated_batches > 32): concatenate neighbouring leaves so that at most
    32 chunks reach the pack kernel.  The reference's cat_input_tensor accepts any count."""
    if len(ts) <= MAX_CHUNKS:
        return ts
    per = -(-len(ts) // MAX_CHUNKS)
    return [torch.cat(ts[i:i + per], 0) if len(ts[i:i + per]) > 1 else ts[i] for i in range(0, len(ts), per)]


def clip_loss_raw(xs, ys, compute: int, want_grad: bool, want_lse: bool = False):
    """xs / ys: lists of [rows_i, D] CUDA tensors (any strides).  Returns (loss[1], dx, dy, row_lse, col_lse) with
    dx / dy packed [B, D] fp32 for grad_output == 1 (None unless want_grad)."""
    L.require_cuda(*xs, *ys)
    xs, ys = _limit_chunks(_as_rows(xs)), _limit_chunks(_as_rows(ys))
    if ys[0].dtype != xs[0].dtype:
        ys = [t.to(xs[0].dtype) for t in ys]
    dev = xs[0].device
    B = sum(t.shape[0] for t in xs)
    D = xs[0].shape[1]
    if sum(t.shape[0] for t in ys) != B or any(t.shape[1] != D for t in xs + ys):
        raise L.PclipError("clip_loss: x and y must hold the same number of [*, D] rows")
    lib = L.lib()
    if not want_lse and _small_fused_ok(B, D, len(xs), len(ys)):
        # small batch (the GradCache configuration): ONE cooperative launch instead of pack / sum-exp / finalize / grad
        loss = torch.empty(1, dtype=torch.float32, device=dev)
        dxy = torch.empty(2, B, D, dtype=torch.float32, device=dev) if want_grad else None
        _small_fused(xs, ys, B, D, None, loss, dxy)
        return loss, (dxy[0] if want_grad else None), (dxy[1] if want_grad else None), None, None
    with torch.cuda.device(dev):     # the workspace plan depends on the SM count of the tensors' device
        nbytes = lib.pclip_clip_workspace_bytes(B, D, compute)
    owner, ws = L.workspace(nbytes, dev)
    loss = torch.empty(1, dtype=torch.float32, device=dev)
    # one allocation for both gradients: backward scales it by grad_output with a single launch
    dxy = torch.empty(2, B, D, dtype=torch.float32, device=dev) if want_grad else None
    dx, dy = (dxy[0], dxy[1]) if want_grad else (None, None)
    rl = torch.empty(B, dtype=torch.float32, device=dev) if want_lse else None
    cl = torch.empty(B, dtype=torch.float32, device=dev) if want_lse else None
    with torch.cuda.device(dev):
        L.check(lib.pclip_clip_loss(L.chunks_of(xs), len(xs), L.chunks_of(ys), len(ys), L.dtype_code(xs[0].dtype), B, D,
                                    compute, L.ptr(loss), L.ptr(rl), L.ptr(cl), L.ptr(dx), L.ptr(dy), ws, nbytes,
                                    L.stream_ptr(dev)))
    owner.record_stream(torch.cuda.current_stream(dev))
    return loss, dx, dy, rl, cl


def _small_fused_ok(B: int, D: int, nx: int, ny: int) -> bool:
    return (D == 128 and B <= int(L.lib().pclip_clip_small_max_batch()) and nx <= MAX_CHUNKS and ny <= MAX_CHUNKS
            and os.environ.get("PCLIP_SMALL_FUSED", "1") != "0")


_small_ws = {}      # device -> (owner, aligned pointer, bytes): the fused kernel's 3 B floats of scratch, stream-ordered


def _small_workspace(dev, B):
    ent = _small_ws.get(dev)
    need = int(L.lib().pclip_clip_small_workspace_bytes(int(L.lib().pclip_clip_small_max_batch())))
    if ent is None:
        owner, p = L.workspace(need, dev)
        ent = _small_ws[dev] = (owner, p, need)
    return ent[1], ent[2]


def _out_chunks(buf_ptr: int, row_counts, D: int):
    """Chunk descriptors of consecutive row blocks of one contiguous fp32 [sum(rows), D] buffer (no tensor objects)."""
    import struct
    flat, r = [], 0
    for n in row_counts:
        flat += (buf_ptr + r * D * 4, n, D, 1)
        r += n
    return (L.Chunk * len(row_counts)).from_buffer_copy(struct.pack(f"{len(flat)}q", *flat))


def _small_fused(xs, ys, B, D, gscale, loss, grad_buf):
    """pclip_clip_small_fused on row chunks xs / ys.  grad_buf: contiguous fp32 [2 B, D] (rows of x, then rows of y)
    that receives the gradients chunk by chunk, or None for the loss only.  Kernels on one stream run in order, so the
    scratch is one persistent buffer per device."""
    lib = L.lib()
    dev = xs[0].device
    with torch.cuda.device(dev):
        ws, nbytes = _small_workspace(dev, B)
        gx = gy = None
        if grad_buf is not None:
            base = grad_buf.data_ptr()
            gx = _out_chunks(base, [t.shape[0] for t in xs], D)
            gy = _out_chunks(base + B * D * 4, [t.shape[0] for t in ys], D)
        L.check(lib.pclip_clip_small_fused(L.chunks_of(xs), len(xs), L.chunks_of(ys), len(ys), L.dtype_code(xs[0].dtype),
                                           B, D, L.ptr(gscale), L.ptr(loss), gx, gy, ws, nbytes, L.stream_ptr(dev)))


def clip_loss_gc_direct(xs, ys, gscale: torch.Tensor | None = None) -> torch.Tensor:
    """GradCache fast path (ref: modules/training_utils.py:137-139: loss = _contrastive_loss_gc(cache_x, cache_y);
    scaler.scale(loss).backward()): ONE kernel launch computes the loss and writes d(gscale * loss) / d(leaf) into the
    `.grad` of every leaf -- no autograd graph over the 32 leaves, no host sync (gscale is GradScaler's device-side
    scale).  Falls back to the autograd path when the batch does not fit the fused kernel.  Returns the (unscaled) loss."""
    xs, ys = list(xs), list(ys)
    L.require_cuda(*xs, *ys)
    B, D = sum(t.shape[0] for t in xs), xs[0].shape[1]
    same = all(t.dtype == xs[0].dtype for t in xs + ys) and xs[0].dtype in (torch.float32, torch.bfloat16)
    if not (same and sum(t.shape[0] for t in ys) == B and _small_fused_ok(B, D, len(xs), len(ys))):
        loss = clip_loss(xs, ys, compute=BF16 if D == 128 else F32)
        (loss if gscale is None else loss * gscale.reshape(())).backward()
        return loss.detach()
    dev = xs[0].device
    loss = torch.empty(1, dtype=torch.float32, device=dev)
    buf = torch.empty(2 * B, D, dtype=torch.float32, device=dev)
    gs = None if gscale is None else gscale.detach().reshape(1).float()
    _small_fused(xs, ys, B, D, gs, loss, buf)
    for leaf, g in zip(xs + ys, buf.split([t.shape[0] for t in xs + ys])):
        if leaf.requires_grad:
            g = g if leaf.dtype == torch.float32 else g.to(leaf.dtype)
            leaf.grad = g if leaf.grad is None else leaf.grad + g
    return loss.reshape(())


class _ClipInfoNCE(torch.autograd.Function):
    """loss = ClipInfoNCE(nx, compute, *x_chunks, *y_chunks).  Gradients are produced together with the loss (one
    fused pass over the score tiles) and scaled by grad_output on the device in backward, so GradScaler's 65536
    (ref: modules/training_utils.py:139) never needs a host sync."""

    @staticmethod
    def forward(ctx, nx, compute, *tensors):
        xs, ys = list(tensors[:nx]), list(tensors[nx:])
        need = any(t.requires_grad for t in tensors)
        # grad mode is off inside forward and only the data pointers are read: no per-leaf detach
        loss, dx, dy, _, _ = clip_loss_raw(xs, ys, compute, need)
        ctx.nx = nx
        ctx.rows = [t.shape[0] for t in tensors]
        ctx.dtypes = [t.dtype for t in tensors]
        if need:
            ctx.save_for_backward(dx._base)         # the [2, B, D] buffer dx and dy are views of
        return loss.reshape(())

    @staticmethod
    def backward(ctx, g):
        (dxy,) = ctx.saved_tensors                      # [2, B, D] for grad_output == 1
        g = g.reshape(1).float().contiguous()
        out = torch.empty_like(dxy)
        with torch.cuda.device(dxy.device):
            L.check(L.lib().pclip_scale_by_scalar(L.ptr(dxy), L.ptr(g), dxy.numel(), L.ptr(out),
                                                  L.stream_ptr(dxy.device)))
        # the 2B rows in leaf order: x chunks then y chunks
        pieces = out.view(-1, out.shape[-1]).split(ctx.rows)
        grads = [p if dt == torch.float32 else p.to(dt) for p, dt in zip(pieces, ctx.dtypes)]
        return (None, None, *grads)


def clip_loss(x, y, compute: int | None = None) -> torch.Tensor:
    """Symmetric InfoNCE of S = x @ y.T.  x / y: [B, D] tensors (any strides) or lists of [rows_i, D] tensors."""
    xs = list(x) if isinstance(x, (list, tuple)) else [x]
    ys = list(y) if isinstance(y, (list, tuple)) else [y]
    if compute is None:
        compute = default_compute()
    return _ClipInfoNCE.apply(len(xs), compute, *xs, *ys)


# ---------------------------------------------------------------------------------------------------------
# Staged CLIP step with caller-owned buffers: pack -> row/column sum-exp -> finalize -> gradients.
# The same stages pclip_clip_loss runs internally, exposed one by one so that (a) a row-sharded job can put
# its all-gather between `pack` and `rowsumexp` (pclip/sharded.py) and (b) bench.py can bracket the dominant
# kernel with CUDA events inside the timed region.  No allocation happens after construction.
# ---------------------------------------------------------------------------------------------------------
class ClipPlan:
    """Buffers and stage calls for S = A Bfull^T with `Ma` local rows per modality against `Nb` gathered rows.

    Single GPU: Ma == Nb == B, diag_offset == 0.  Row-sharded: Ma = B / P, Nb = B, diag_offset = rank * Ma."""

    def __init__(self, Ma: int, Nb: int, D: int, compute: int, device, diag_offset: int = 0):
        self.Ma, self.Nb, self.D, self.compute, self.dev, self.off = Ma, Nb, D, compute, torch.device(device), diag_offset
        tc = compute == BF16
        self.cdt = torch.bfloat16 if tc else torch.float32
        self.ldT = (Nb + 63) // 64 * 64
        dev = self.dev
        f32 = dict(dtype=torch.float32, device=dev)
        # local packed rows of both modalities, and the full (gathered) matrices; on one GPU they alias
        self.xl = torch.empty(Ma, D, dtype=self.cdt, device=dev)
        self.yl = torch.empty(Ma, D, dtype=self.cdt, device=dev)
        self.local_is_full = Ma == Nb
        self.xf = self.xl if self.local_is_full else torch.empty(Nb, D, dtype=self.cdt, device=dev)
        self.yf = self.yl if self.local_is_full else torch.empty(Nb, D, dtype=self.cdt, device=dev)
        # second-generation bf16 kernels: single-pass forward (row AND column sums from one sweep over S) and a
        # gradient kernel that needs no transposed operand copies
        # (below SINGLE_PASS_MIN_COLS columns the two-sided forward is used: its column sums are fp32-exact, whereas the
        # single pass sums bf16-rounded exps -- errors that average out over thousands of rows but not over 16)
        self.single_pass = tc and Nb >= SINGLE_PASS_MIN_COLS
        self.xfT = self.yfT = None      # no kernel reads transposed operand copies any more (fields kept in the ABI)
        self.rowsum, self.colsum = torch.empty(Ma, **f32), torch.empty(Ma, **f32)
        # single-pass: column sums of ALL Nb columns over this plan's rows (partial under row sharding)
        self.colsum_f = self.colsum if self.local_is_full else torch.empty(Nb, **f32)
        self.diag, self.diag2 = torch.empty(Ma, **f32), torch.empty(Ma, **f32)
        self.row_lse, self.col_lse = torch.empty(Ma, **f32), torch.empty(Ma, **f32)
        self.rinv, self.cinv = torch.empty(Ma, **f32), torch.empty(Ma, **f32)
        # 1/R and 1/C of ALL rows (the gradient kernel needs the other side's normalisers for every column)
        self.rinv_f = self.rinv if self.local_is_full else torch.empty(Nb, **f32)
        self.cinv_f = self.cinv if self.local_is_full else torch.empty(Nb, **f32)
        if self.single_pass:
            # What the forward sweep produces is ONE contiguous record [C partial (Nb) | R (Ma) | S_ii (Ma)]: a row shard
            # owes it to the other ranks (pclip/sharded.py gathers it in one collective), and pclip_clip_finalize_sharded
            # turns P such records (P = 1 on one GPU: the same kernel, the same summation order, the same loss bits)
            # into the loss and 1/R, 1/C of ALL rows; the local 1/R, 1/C are views of the full-length results.
            self.stats_l = torch.zeros(Nb + 2 * Ma, **f32)
            self.colsum_f = self.stats_l[:Nb]
            self.rowsum = self.stats_l[Nb:Nb + Ma]
            self.diag = self.stats_l[Nb + Ma:]
            self.row_lse_f, self.col_lse_f = torch.empty(Nb, **f32), torch.empty(Nb, **f32)
            self.rinv_f, self.cinv_f = torch.empty(Nb, **f32), torch.empty(Nb, **f32)
            self.rinv = self.rinv_f[diag_offset:diag_offset + Ma]
            self.cinv = self.cinv_f[diag_offset:diag_offset + Ma]
            self.row_lse = self.row_lse_f[diag_offset:diag_offset + Ma]
            self.col_lse = self.col_lse_f[diag_offset:diag_offset + Ma]
            if self.local_is_full:
                self.colsum = self.colsum_f
            with torch.cuda.device(dev):
                self.fin_bytes = int(L.lib().pclip_clip_finalize_sharded_workspace_bytes(Nb))
            self._fin_owner, self.fin_ws = L.workspace(self.fin_bytes, dev)
        self.loss = torch.zeros(1, **f32)
        self.dx, self.dy = torch.empty(Ma, D, **f32), torch.empty(Ma, D, **f32)
        lib = L.lib()
        with torch.cuda.device(dev):
            n = max(lib.pclip_clip_rowsumexp_workspace_bytes(Ma, Nb, D, compute),
                    lib.pclip_clip_sumexp_workspace_bytes(Ma, Nb, D, compute) if tc else 0,
                    lib.pclip_clip_grad_workspace_bytes(Ma, Nb, D, compute), 1024)
        self.ws_bytes = int(n)
        self._ws_owner, self.ws = L.workspace(n, dev)

    def _st(self):
        return L.stream_ptr(self.dev)

    def _on_device(fn):   # every stage launches on the plan's device, whatever the caller's current device is
        def wrapped(self, *a, **k):
            with torch.cuda.device(self.dev):
                return fn(self, *a, **k)
        wrapped.__name__, wrapped.__doc__ = fn.__name__, fn.__doc__
        return wrapped

    @_on_device
    def pack(self, xs, ys):
        """Strided row chunks (GradCache leaves, `.t()` views) -> packed compute-dtype rows (+ transposes)."""
        lib = L.lib()
        xs, ys = _limit_chunks(_as_rows(xs)), _limit_chunks(_as_rows(ys))
        full = self.local_is_full
        if self.xfT is None and xs[0].dtype == ys[0].dtype:      # both modalities, one launch
            L.check(lib.pclip_pack_pair(L.chunks_of(xs), len(xs), L.chunks_of(ys), len(ys), L.dtype_code(xs[0].dtype),
                                        self.Ma, self.D, L.dtype_code(self.cdt), L.ptr(self.xl), L.ptr(self.yl),
                                        self._st()))
            return
        for src, out, outT in ((xs, self.xl, self.xfT if full else None), (ys, self.yl, self.yfT if full else None)):
            L.check(lib.pclip_pack(L.chunks_of(src), len(src), L.dtype_code(src[0].dtype), self.Ma, self.D,
                                   L.dtype_code(self.cdt), L.ptr(out), L.ptr(outT), self.ldT, self._st()))

    @_on_device
    def transpose_full(self):
        """After an all-gather into xf / yf: the transposed copies the bf16 gradient kernel reads."""
        if self.xfT is None or self.local_is_full:
            return
        lib = L.lib()
        for src, dst in ((self.xf, self.xfT), (self.yf, self.yfT)):
            L.check(lib.pclip_transpose_bf16(L.ptr(src), self.Nb, self.D, L.ptr(dst), self.ldT, self._st()))

    @_on_device
    def rowsumexp(self):
        sides = (L.FwdSide * 2)(
            L.FwdSide(L.ptr(self.xl), L.ptr(self.yf), self.Ma, self.Nb, self.off, L.ptr(self.rowsum), L.ptr(self.diag)),
            L.FwdSide(L.ptr(self.yl), L.ptr(self.xf), self.Ma, self.Nb, self.off, L.ptr(self.colsum), L.ptr(self.diag2)))
        L.check(L.lib().pclip_clip_rowsumexp(sides, 2, self.D, self.compute, self.ws, self.ws_bytes, self._st()))

    @_on_device
    def sumexp(self):
        """Single-pass forward: rowsum (complete), colsum_f (all Nb columns, summed over THIS plan's rows), diag."""
        L.check(L.lib().pclip_clip_sumexp(L.ptr(self.xl), L.ptr(self.yf), self.Ma, self.Nb, self.off, self.D,
                                          self.compute, L.ptr(self.rowsum), L.ptr(self.colsum_f), L.ptr(self.diag),
                                          self.ws, self.ws_bytes, self._st()))

    def take_local_colsum(self):
        """After the all-reduce of colsum_f over ranks: the column sums of this rank's own positives."""
        if not self.local_is_full:
            self.colsum.copy_(self.colsum_f[self.off:self.off + self.Ma])

    @_on_device
    def finalize(self, B_total: int):
        """loss[0] = 1/(2B) sum_i (log R_i + log C_i - 2 S_ii) over the local rows; 1/R, 1/C for the backward."""
        if self.single_pass and self.local_is_full:
            return self.finalize_full(self.stats_l.view(1, -1), B_total)
        L.check(L.lib().pclip_clip_finalize(L.ptr(self.rowsum), L.ptr(self.colsum), L.ptr(self.diag), self.Ma,
                                            0.5 / B_total, L.ptr(self.row_lse), L.ptr(self.col_lse), L.ptr(self.rinv),
                                            L.ptr(self.cinv), L.ptr(self.loss), self._st()))

    @_on_device
    def finalize_full(self, stats_all: torch.Tensor, B_total: int):
        """Row shard, after the all-gather of every rank's [C partial | R | S_ii] record (`stats_all` [P, Nb + 2 Ma]):
        reduce the column partials, lay R and S_ii out by global row, and finalise ALL rows here -- the loss comes out
        global on every rank and 1/R, 1/C of every row are local, so no further exchange is needed."""
        L.check(L.lib().pclip_clip_finalize_sharded(L.ptr(stats_all), stats_all.shape[0], self.Nb, self.Ma,
                                                    0.5 / B_total, L.ptr(self.row_lse_f), L.ptr(self.col_lse_f),
                                                    L.ptr(self.rinv_f), L.ptr(self.cinv_f), L.ptr(self.loss),
                                                    self.fin_ws, self.fin_bytes, self._st()))

    @_on_device
    def grad(self, B_total: int, gscale: torch.Tensor | None = None):
        sides = (L.BwdSide * 2)(
            L.BwdSide(L.ptr(self.xl), L.ptr(self.yf), L.ptr(self.yfT), self.ldT, self.Ma, self.Nb, self.off,
                      L.ptr(self.rinv), L.ptr(self.cinv_f), L.ptr(self.dx)),
            L.BwdSide(L.ptr(self.yl), L.ptr(self.xf), L.ptr(self.xfT), self.ldT, self.Ma, self.Nb, self.off,
                      L.ptr(self.cinv), L.ptr(self.rinv_f), L.ptr(self.dy)))
        L.check(L.lib().pclip_clip_grad(sides, 2, self.D, self.compute, 0.5 / B_total, L.ptr(gscale), self.ws,
                                        self.ws_bytes, self._st()))

    def step(self, xs, ys, events=None):
        """Single-GPU fwd + bwd.  `events`: optional dict name -> (start, end) torch.cuda.Event pairs to record
        around the stages on the current stream (bench.py's roofline legs)."""
        def rec(name, i):
            if events is not None and name in events:
                events[name][i].record(torch.cuda.current_stream(self.dev))
        assert self.local_is_full
        rec("pack", 0); self.pack(xs, ys); rec("pack", 1)
        rec("rowsumexp", 0)
        self.sumexp() if self.single_pass else self.rowsumexp()
        rec("rowsumexp", 1)
        self.finalize(self.Nb)
        rec("grad", 0); self.grad(self.Nb); rec("grad", 1)
        return self.loss, self.dx, self.dy


# ---------------------------------------------------------------------------------------------------------
# (3) scores / rank of the positive / top-k      ref: modules/training_utils.py:157-172, :219-222;
#                                                      modules/visualizations.py:51-52; topk.py:47-61
# ---------------------------------------------------------------------------------------------------------
def _f32c(t):
    return t.detach().float().contiguous()


def scores(q: torch.Tensor, bank: torch.Tensor, op: str = "dot", neg_tau: torch.Tensor | None = None) -> torch.Tensor:
    """[Q, N] matrix of q . bank_n ("dot"), exp(q . bank_n) ("exp": get_logits) and, with neg_tau, times exp(-tau)
    (the cosine matrix of visualizations.py:51-52)."""
    L.require_cuda(q, bank)
    q, bank = _f32c(q), _f32c(bank)
    out = torch.empty(q.shape[0], bank.shape[0], dtype=torch.float32, device=q.device)
    nt = None if neg_tau is None else neg_tau.detach().reshape(1).float().to(q.device)
    with torch.cuda.device(q.device):
        L.check(L.lib().pclip_scores(L.ptr(q), L.ptr(bank), q.shape[0], bank.shape[0], q.shape[1],
                                     L.SCORE_EXP if op == "exp" else L.SCORE_DOT, L.ptr(nt), L.ptr(out),
                                     L.stream_ptr(q.device)))
    return out


def rank_of_positive(q: torch.Tensor, bank: torch.Tensor, pos: torch.Tensor) -> torch.Tensor:
    """rank[i] = 1 + #{n : q_i . bank_n > q_i . bank_pos[i]} (strict), int32 [Q]."""
    L.require_cuda(q, bank, pos)
    q, bank = _f32c(q), _f32c(bank)
    pos = pos.to(torch.int64).contiguous()
    rank = torch.empty(q.shape[0], dtype=torch.int32, device=q.device)
    with torch.cuda.device(q.device):
        nbytes = int(L.lib().pclip_rank_workspace_bytes(q.shape[0]))
        owner, ws = L.workspace(nbytes, q.device)
        L.check(L.lib().pclip_rank_of_positive(L.ptr(q), L.ptr(bank), L.ptr(pos), q.shape[0], bank.shape[0], q.shape[1],
                                               L.ptr(rank), ws, nbytes, L.stream_ptr(q.device)))
    owner.record_stream(torch.cuda.current_stream(q.device))
    return rank


class TopkPlan:
    """Caller-owned outputs and workspace for repeated top-k calls of one shape (no allocation per call)."""

    def __init__(self, Q: int, N: int, D: int, k: int, compute: int, device):
        self.Q, self.N, self.D, self.k, self.compute, self.dev = Q, N, D, k, compute, torch.device(device)
        self.idx = torch.empty(Q, k, dtype=torch.int64, device=device)
        self.score = torch.empty(Q, k, dtype=torch.float32, device=device)
        self.n_unsafe = torch.zeros(1, dtype=torch.int32, device=device)
        with torch.cuda.device(self.dev):
            self.ws_bytes = int(L.lib().pclip_topk_workspace_bytes(Q, N, D, k, compute))
        self._owner, self.ws = L.workspace(max(self.ws_bytes, 1024), device)
        self._bank_key, self._bank_owner, self._bank_img = None, None, None

    def prepare_bank(self, bank: torch.Tensor):
        """Build the persistent bf16 image of `bank` once; later run() calls with the SAME tensor (same storage, not
        modified in place since) skip the per-call fp32 -> bf16 conversion of the bank (3 N D bytes of HBM traffic)."""
        assert self.compute == BF16 and bank.shape == (self.N, self.D) and bank.dtype == torch.float32
        assert bank.is_contiguous()
        lib = L.lib()
        with torch.cuda.device(self.dev):
            nbytes = int(lib.pclip_topk_bank_bytes(self.N, self.D))
            self._bank_owner, self._bank_img = L.workspace(nbytes, self.dev)
            L.check(lib.pclip_topk_prepare_bank(L.ptr(bank), self.N, self.D, self._bank_img, nbytes,
                                                L.stream_ptr(self.dev)))
        self._bank_key = (bank.data_ptr(), bank._version)
        return self

    def run(self, q: torch.Tensor, bank: torch.Tensor, index_base: int = 0):
        assert q.shape == (self.Q, self.D) and bank.shape == (self.N, self.D)
        assert q.dtype == torch.float32 and bank.dtype == torch.float32 and q.is_contiguous() and bank.is_contiguous()
        lib = L.lib()
        with torch.cuda.device(self.dev):
            if self._bank_key == (bank.data_ptr(), bank._version):
                L.check(lib.pclip_topk_prepared(L.ptr(q), L.ptr(bank), self._bank_img, self.Q, self.N, self.D, self.k,
                                                index_base, L.ptr(self.idx), L.ptr(self.score), L.ptr(self.n_unsafe),
                                                self.ws, self.ws_bytes, L.stream_ptr(self.dev)))
            else:
                L.check(lib.pclip_topk(L.ptr(q), L.ptr(bank), self.Q, self.N, self.D, self.k, self.compute, index_base,
                                       L.ptr(self.idx), L.ptr(self.score), L.ptr(self.n_unsafe), self.ws, self.ws_bytes,
                                       L.stream_ptr(self.dev)))
        return self.idx, self.score


def topk(q: torch.Tensor, bank: torch.Tensor, k: int, compute: int | None = None, index_base: int = 0):
    """Top-k bank rows per query by dot product: (idx [Q, k] int64, score [Q, k] fp32), descending, ties -> lower
    index.  BF16 = tensor-core candidate selection + exact fp32 re-scoring (proven exact per query, else recomputed
    by the fp32 kernel); F32 = exact fp32 scan."""
    L.require_cuda(q, bank)
    q, bank = _f32c(q), _f32c(bank)
    if compute is None:
        compute = BF16 if q.shape[1] == 128 else F32
    plan = TopkPlan(q.shape[0], bank.shape[0], q.shape[1], k, compute, q.device)
    idx, score = plan.run(q, bank, index_base)
    plan._owner.record_stream(torch.cuda.current_stream(q.device))
    return idx, score, plan.n_unsafe


def topk_merge(idx_lists: torch.Tensor, score_lists: torch.Tensor):
    """Merge [n_lists, Q, k] per-shard candidates (global indices) into the global top-k (bank-sharded mode)."""
    L.require_cuda(idx_lists, score_lists)
    n, Q, k = idx_lists.shape
    idx = torch.empty(Q, k, dtype=torch.int64, device=idx_lists.device)
    score = torch.empty(Q, k, dtype=torch.float32, device=idx_lists.device)
    with torch.cuda.device(idx.device):
        L.check(L.lib().pclip_topk_merge(L.ptr(idx_lists.contiguous()), L.ptr(score_lists.contiguous()), Q, n, k,
                                         L.ptr(idx), L.ptr(score), L.stream_ptr(idx.device)))
    return idx, score


# ---------------------------------------------------------------------------------------------------------
# (4) FILIP token-wise late interaction    ref: modules/models.py:292-356; modules/training_utils_filip.py:4-39
# ---------------------------------------------------------------------------------------------------------
def mask_lengths(mask: torch.Tensor) -> torch.Tensor:
    """Right-padded attention mask [B, T] (a contiguous prefix of ones, as the ESM tokenizer produces with
    padding=True, ref: modules/training_utils_filip.py:56-57) -> int32 lengths.  Set PCLIP_CHECK_MASKS=1 to verify the
    prefix property (costs a host sync)."""
    m = mask.to(torch.bool)
    lens = m.sum(dim=1).to(torch.int32)
    if os.environ.get("PCLIP_CHECK_MASKS") == "1":
        T = m.shape[1]
        want = torch.arange(T, device=m.device)[None, :] < lens[:, None]
        if not bool((want == m).all()):
            raise L.PclipError("FILIP kernels need right-padded masks (a prefix of ones per row)")
    return lens.contiguous()


def compact_tokens(h: torch.Tensor, mask: torch.Tensor):
    """Move the valid tokens of every sample to the front, keeping their order: ([B, T, D] reordered, int32 lengths).
    FILIP's max-over-tokens and masked mean do not depend on token order, so an arbitrary 0/1 mask (the reference
    accepts any, modules/models.py:333-340) becomes the right-padded form the kernels take; pads end up behind the
    valid tokens with their values untouched (they are never read).  Differentiable through the gather; no host sync."""
    m = mask.to(torch.bool)
    order = torch.sort((~m).to(torch.int8), dim=1, stable=True).indices          # valid tokens first, stable
    lens = m.sum(dim=1).to(torch.int32)
    return torch.gather(h, 1, order[..., None].expand_as(h)), lens


class _FilipSimilarity(torch.autograd.Function):
    @staticmethod
    def forward(ctx, hA, hB, lenA, lenB, tau, compute):
        L.require_cuda(hA, hB, lenA, lenB)
        dev = hA.device
        a, b = hA.detach().float().contiguous(), hB.detach().float().contiguous()
        Ba, TA, D = a.shape
        Bb, TB, _ = b.shape
        tau_d = tau.detach().reshape(1).float().to(dev)
        lib = L.lib()
        with torch.cuda.device(dev):
            nbytes = int(lib.pclip_filip_workspace_bytes(Ba, Bb, TA, TB, D, compute))
        owner, ws = L.workspace(nbytes, dev)
        simA = torch.empty(Ba, Bb, dtype=torch.float32, device=dev)
        simB = torch.empty(Ba, Bb, dtype=torch.float32, device=dev)
        argA = torch.empty(Ba, Bb, TA, dtype=torch.int16, device=dev)
        argB = torch.empty(Bb, Ba, TB, dtype=torch.int16, device=dev)
        with torch.cuda.device(dev):
            L.check(lib.pclip_filip_fwd(L.ptr(a), L.ptr(b), L.ptr(lenA), L.ptr(lenB), Ba, Bb, TA, TB, D, L.ptr(tau_d),
                                        compute, L.ptr(simA), L.ptr(simB), L.ptr(argA), L.ptr(argB), ws, nbytes,
                                        L.stream_ptr(dev)))
        owner.record_stream(torch.cuda.current_stream(dev))
        ctx.save_for_backward(a, b, lenA, lenB, tau_d, simA, simB, argA, argB)
        ctx.compute = compute
        ctx.tau_shape = tau.shape
        ctx.in_dtypes = (hA.dtype, hB.dtype)
        return simA, simB

    @staticmethod
    def backward(ctx, gA, gB):
        a, b, lenA, lenB, tau_d, simA, simB, argA, argB = ctx.saved_tensors
        dev = a.device
        Ba, TA, D = a.shape
        Bb, TB, _ = b.shape
        gA = torch.zeros_like(simA) if gA is None else gA.float().contiguous()
        gB = torch.zeros_like(simB) if gB is None else gB.float().contiguous()
        lib = L.lib()
        with torch.cuda.device(dev):
            nbytes = int(lib.pclip_filip_workspace_bytes(Ba, Bb, TA, TB, D, ctx.compute))
        owner, ws = L.workspace(nbytes, dev)
        dhA, dhB = torch.empty_like(a), torch.empty_like(b)
        dtau = torch.zeros(1, dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            L.check(lib.pclip_filip_bwd(L.ptr(a), L.ptr(b), L.ptr(lenA), L.ptr(lenB), Ba, Bb, TA, TB, D, L.ptr(tau_d),
                                        ctx.compute, L.ptr(simA), L.ptr(simB), L.ptr(argA), L.ptr(argB), L.ptr(gA),
                                        L.ptr(gB), L.ptr(dhA), L.ptr(dhB), L.ptr(dtau), ws, nbytes, L.stream_ptr(dev)))
        owner.record_stream(torch.cuda.current_stream(dev))
        return (dhA.to(ctx.in_dtypes[0]), dhB.to(ctx.in_dtypes[1]), None, None, dtau.reshape(ctx.tau_shape), None)


def filip_similarity_raw(hA, hB, lenA, lenB, tau, compute: int):
    """pclip_filip_fwd without autograd: (simA, simB [Ba,Bb], argA [Ba,Bb,TA], argB [Bb,Ba,TB] int16) -- the stored
    maximisers are what the backward consumes; parity tests compare them between compute modes."""
    L.require_cuda(hA, hB, lenA, lenB)
    dev = hA.device
    a, b = hA.detach().float().contiguous(), hB.detach().float().contiguous()
    Ba, TA, D = a.shape
    Bb, TB, _ = b.shape
    tau_d = torch.as_tensor(tau, device=dev).detach().reshape(1).float()
    la, lb = lenA.to(torch.int32).contiguous(), lenB.to(torch.int32).contiguous()
    lib = L.lib()
    simA = torch.empty(Ba, Bb, dtype=torch.float32, device=dev)
    simB = torch.empty(Ba, Bb, dtype=torch.float32, device=dev)
    argA = torch.zeros(Ba, Bb, TA, dtype=torch.int16, device=dev)
    argB = torch.zeros(Bb, Ba, TB, dtype=torch.int16, device=dev)
    with torch.cuda.device(dev):
        nbytes = int(lib.pclip_filip_workspace_bytes(Ba, Bb, TA, TB, D, compute))
        owner, ws = L.workspace(nbytes, dev)
        L.check(lib.pclip_filip_fwd(L.ptr(a), L.ptr(b), L.ptr(la), L.ptr(lb), Ba, Bb, TA, TB, D, L.ptr(tau_d), compute, L.ptr(simA), L.ptr(simB), L.ptr(argA), L.ptr(argB),
                                    ws, nbytes, L.stream_ptr(dev)))
    owner.record_stream(torch.cuda.current_stream(dev))
    return simA, simB, argA, argB


def filip_similarity(hA, hB, maskA, maskB, tau, compute: int | None = None):
    """(simA, simB) [Ba, Bb] of ExtendedFILIP._filip_similarity_score for right-padded masks (or int32 lengths)."""
    if compute is None:
        compute = default_compute()
    lenA = maskA if maskA.dim() == 1 else mask_lengths(maskA)
    lenB = maskB if maskB.dim() == 1 else mask_lengths(maskB)
    tau = torch.as_tensor(tau, device=hA.device)
    return _FilipSimilarity.apply(hA, hB, lenA.to(torch.int32), lenB.to(torch.int32), tau, compute)


class _RowwiseInfoNCE(torch.autograd.Function):
    """mean_i(-L_ii + log sum_j exp L_ij) * weight  (ref: modules/training_utils_filip.py:15-23, use_dcl=False)."""

    @staticmethod
    def forward(ctx, logits, weight, diag_offset):
        L.require_cuda(logits)
        lg = logits.detach().float().contiguous()
        B, Bc = lg.shape
        loss = torch.zeros(1, dtype=torch.float32, device=lg.device)
        dl = torch.empty_like(lg) if logits.requires_grad else None
        with torch.cuda.device(lg.device):
            nbytes = int(L.lib().pclip_rowsum_workspace_bytes(B))
            owner, ws = L.workspace(nbytes, lg.device)
            L.check(L.lib().pclip_rowwise_infonce(L.ptr(lg), B, Bc, diag_offset, float(weight) / B, L.ptr(loss), L.ptr(dl),
                                                  ws, nbytes, L.stream_ptr(lg.device)))
        owner.record_stream(torch.cuda.current_stream(lg.device))
        if dl is not None:
            ctx.save_for_backward(dl)
        ctx.in_dtype = logits.dtype
        return loss.reshape(())

    @staticmethod
    def backward(ctx, g):
        (dl,) = ctx.saved_tensors
        g = g.reshape(1).float().contiguous()
        out = torch.empty_like(dl)
        with torch.cuda.device(dl.device):
            L.check(L.lib().pclip_scale_by_scalar(L.ptr(dl), L.ptr(g), dl.numel(), L.ptr(out), L.stream_ptr(dl.device)))
        return out.to(ctx.in_dtype), None, None


def rowwise_infonce(logits: torch.Tensor, weight: float = 1.0, diag_offset: int = 0) -> torch.Tensor:
    return _RowwiseInfoNCE.apply(logits, weight, diag_offset)


def filip_loss(simA: torch.Tensor, simB: torch.Tensor, diag_offset: int = 0) -> torch.Tensor:
    """(loss_A + loss_B) / 2 with BOTH matrices normalised row-wise (ref: training_utils_filip.py:26-39)."""
    return rowwise_infonce(simA, 0.5, diag_offset) + rowwise_infonce(simB, 0.5, diag_offset)


# ---------------------------------------------------------------------------------------------------------
# (5) fused encoder head, forward      ref: modules/models.py:46-51 and :192-196
# ---------------------------------------------------------------------------------------------------------
def _ffn_layers(ffn):
    """nn.Sequential built by _build_ffn (ref: modules/models.py:66-73) -> list of (Linear, LayerNorm | None), or None
    when the module does not have that shape (then the caller keeps the stock path)."""
    out = []
    mods = list(ffn)
    for i, m in enumerate(mods):
        if isinstance(m, torch.nn.Linear) and i == len(mods) - 1:
            out.append((m, None))
        elif isinstance(m, torch.nn.Sequential) and len(m) == 4 and isinstance(m[0], torch.nn.Linear) and \
                isinstance(m[1], torch.nn.ReLU) and isinstance(m[2], torch.nn.LayerNorm) and \
                isinstance(m[3], torch.nn.Dropout) and m[2].elementwise_affine and m[2].bias is not None:
            out.append((m[0], m[2]))
        else:
            return None
    return out


def _linear_array(layers, keep):
    arr = (L.Linear * max(len(layers), 1))()
    for i, (lin, ln) in enumerate(layers):
        ts = [lin.weight, lin.bias, ln.weight if ln is not None else None, ln.bias if ln is not None else None]
        ts = [None if t is None else t.detach().float().contiguous() for t in ts]
        keep.extend(ts)
        arr[i] = L.Linear(*[L.ptr(t) for t in ts])
    return arr


def encoder_head_supported(hidden, projection, token_ffn, pooled_ffn) -> bool:
    """True when the fused kernels cover this head: CUDA fp32 parameters, width 128, Din % 4 == 0, FFNs of the reference's
    shape with at most 8 layers, biases present, one LayerNorm eps."""
    if not (hidden.is_cuda and projection.weight.is_cuda and projection.bias is not None):
        return False
    if projection.out_features != 128 or projection.in_features % 4 or hidden.shape[-1] != projection.in_features:
        return False
    eps = set()
    for ffn in (token_ffn, pooled_ffn):
        if ffn is None:
            continue
        layers = _ffn_layers(ffn)
        if layers is None or len(layers) > 8:
            return False
        for lin, ln in layers:
            if lin.in_features != 128 or lin.out_features != 128 or lin.bias is None or lin.weight.dtype != torch.float32:
                return False
            if ln is not None:
                eps.add(ln.eps)
    return len(eps) <= 1 and projection.weight.dtype == torch.float32


def encoder_head(hidden, mask, projection, token_ffn, pooled_ffn=None, tau=None, norm_eps: float = 1e-12):
    """hidden [Bs, T, Din] (the ESM trunk's last_hidden_state), mask [Bs, T].  pooled_ffn None: per-residue embeddings
    [Bs, T, 128] (FILIPEncoder); else F.normalize(pooled_ffn(masked_mean(...))) * exp(tau / 2) [Bs, 128] (Encoder).
    Inference only (no autograd graph is recorded)."""
    L.require_cuda(hidden, mask)
    dev = hidden.device
    hid = hidden.detach().float().contiguous()
    m = mask.detach().to(torch.int64).contiguous()
    Bs, T, Din = hid.shape
    keep = []
    tok_layers = _ffn_layers(token_ffn)
    proj = _linear_array([(projection, None)], keep)
    tok = _linear_array(tok_layers, keep)
    pooled = pooled_ffn is not None
    pl_layers = _ffn_layers(pooled_ffn) if pooled else []
    pl = _linear_array(pl_layers, keep)
    lns = [ln for _, ln in tok_layers + pl_layers if ln is not None]
    ln_eps = float(lns[0].eps) if lns else 1e-5
    out = torch.empty((Bs, 128) if pooled else (Bs, T, 128), dtype=torch.float32, device=dev)
    if Bs == 0:
        return out
    t = None if tau is None else torch.as_tensor(tau).detach().float().reshape(1).to(dev)
    lib = L.lib()
    with torch.cuda.device(dev):
        n = int(lib.pclip_encoder_head_workspace_bytes(Bs, T)) if pooled else 0
        owner, ws = L.workspace(n, dev) if pooled else (None, None)
        L.check(lib.pclip_encoder_head_fwd(L.ptr(hid), L.ptr(m), Bs, T, Din, 128, proj, tok, len(tok_layers), pl,
                                           len(pl_layers), L.ptr(t), ln_eps, float(norm_eps),
                                           None if pooled else L.ptr(out), L.ptr(out) if pooled else None, ws, n,
                                           L.stream_ptr(dev)))
    return out
