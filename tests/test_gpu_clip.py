"""GPU parity of the CLIP InfoNCE path (libpclip.so through the C ABI) against the oracle and the golden vectors
generated from the reference's own functions.  Tolerances are north_star's: 1e-5 relative in fp32, 1e-2 in bf16."""
import math
import os

import pytest
import torch

from oracle import restate as R

pytestmark = pytest.mark.gpu

F64 = torch.float64
TOL = {"f32": 1e-5, "bf16": 1e-2}


def rel_err(a, b):
    a = torch.as_tensor(a).detach().double().cpu()
    b = torch.as_tensor(b).detach().double().cpu()
    assert a.shape == b.shape, (a.shape, b.shape)
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def assert_close(a, b, tol, what, atol=0.0):
    e = rel_err(a, b)
    if atol:
        d = float((torch.as_tensor(a).detach().double().cpu() - torch.as_tensor(b).detach().double().cpu()).abs().max())
        if d <= atol:
            return
    assert e <= tol, f"{what}: max rel err {e:.3e} > {tol:.1e}"


@pytest.fixture(scope="module")
def ops():
    from pclip import ops as o
    return o


def _mode(ops, name):
    return ops.F32 if name == "f32" else ops.BF16


def test_library_loads_and_counts_launches(ops):
    from pclip import _lib
    assert _lib.lib().pclip_version() >= 100
    ops.reset_launch_count()
    x = torch.randn(16, 128, device="cuda")
    ops.clip_loss(x, x)
    assert ops.launch_count() == 1          # B <= 512: the one-launch small-batch kernel
    ops.reset_launch_count()
    big = torch.randn(1024, 128, device="cuda") / 11.3
    ops.clip_loss(big, big, compute=ops.BF16)
    assert ops.launch_count() >= 3         # pack, sum-exp, finalize (no gradient wanted)


def test_encoder_tail(ops, golden):
    g = golden("encoder_tail")
    e = g["e"].cuda().requires_grad_()
    tau = g["tau"].cuda().requires_grad_()
    out = ops.normalize_scale(e, tau)
    assert_close(out, g["out"], 1e-6, "normalize_scale")
    out.backward(g["g"].cuda())
    assert_close(e.grad, g["de"], 1e-5, "de")
    assert_close(tau.grad, g["dtau"], 1e-5, "dtau")


@pytest.mark.parametrize("name", ["clip_b16", "clip_b37", "clip_b256"])
@pytest.mark.parametrize("mode", ["f32", "bf16"])
def test_clip_golden(ops, golden, name, mode):
    """Encoder tail + _contrastive_loss(pep, rec.t()) forward and backward (ref: training_utils.py:68-71)."""
    g = golden(name)
    e_pep = g["e_pep"].cuda().requires_grad_()
    e_rec = g["e_rec"].cuda().requires_grad_()
    tau = g["tau"].cuda().requires_grad_()
    pep = ops.normalize_scale(e_pep, tau)
    rec = ops.normalize_scale(e_rec, tau)
    loss = ops.clip_loss(pep, rec.t().t(), compute=_mode(ops, mode))
    tol = TOL[mode]
    assert_close(loss, g["loss"], tol, "loss")
    loss.backward()
    assert_close(e_pep.grad, g["de_pep"], tol * 3, "de_pep")
    assert_close(e_rec.grad, g["de_rec"], tol * 3, "de_rec")
    assert_close(tau.grad, g["dtau"], tol * 3, "dtau")


@pytest.mark.parametrize("mode", ["f32", "bf16"])
def test_clip_transposed_view_and_trained_like(ops, golden, mode):
    g = golden("clip_trained_like")
    pep = g["pep"].cuda().requires_grad_()
    rec = g["rec"].cuda().requires_grad_()
    recT = rec.t()  # the reference passes rec.t() (stride (1, D)); we undo the view without copying
    loss = ops.clip_loss(pep, recT.t(), compute=_mode(ops, mode))
    loss.backward()
    tol = TOL[mode]
    assert_close(loss, g["loss"], tol, "loss")
    assert_close(pep.grad, g["dpep"], tol * 3, "dpep")
    assert_close(rec.grad, g["drec"], tol * 3, "drec")
    # a genuinely strided operand: columns-major storage
    rec_cm = g["rec"].cuda().t().contiguous().t().requires_grad_()
    assert rec_cm.stride() == (1, rec_cm.shape[0])
    loss2 = ops.clip_loss(g["pep"].cuda(), rec_cm, compute=_mode(ops, mode))
    assert_close(loss2, g["loss"], tol, "loss (column-major y)")


@pytest.mark.parametrize("mode", ["f32", "bf16"])
def test_gradcache_leaves(ops, golden, mode):
    """16 + 16 detached leaves of [16,128], grad_output = 65536 (ref: training_utils.py:137-139, :247-265)."""
    g = golden("clip_gradcache_16x16")
    xs = [t.cuda().requires_grad_() for t in g["x"]]
    ys = [t.cuda().requires_grad_() for t in g["y"]]
    loss = ops.clip_loss(xs, ys, compute=_mode(ops, mode))
    tol = TOL[mode]
    assert_close(loss, g["loss"], tol, "loss")
    (loss * float(g["grad_output"])).backward()
    dx = torch.stack([t.grad for t in xs])
    dy = torch.stack([t.grad for t in ys])
    assert_close(dx, g["dx"], tol * 3, "dx")
    assert_close(dy, g["dy"], tol * 3, "dy")


@pytest.mark.parametrize("B", [1, 5, 129, 200, 1000])
def test_clip_bf16_ragged_sizes_vs_oracle_on_rounded_inputs(ops, B):
    """The tensor-core path against the float64 oracle evaluated on the SAME bf16-rounded operands: isolates kernel
    arithmetic (fp32 accumulate, ex2.approx, bf16 re-quantised P) from input rounding."""
    from pclip import synth
    x, y = synth.clip_scaled_embeddings(B, 128, cfg_id=100 + B)
    xr, yr = R.bf16_round(x), R.bf16_round(y)
    loss_o, dx_o, dy_o, rl_o, cl_o = R.clip_loss_and_grads(xr.double(), yr.double())
    loss, dx, dy, rl, cl = ops.clip_loss_raw([x.cuda()], [y.cuda()], ops.BF16, True, want_lse=True)
    assert_close(loss[0], loss_o, 1e-5, "loss", atol=1e-6)  # B = 1: the loss is exactly 0
    assert_close(rl, rl_o, 1e-5, "row_lse")
    assert_close(cl, cl_o, 1e-5, "col_lse")
    assert_close(dx, dx_o, 4e-3, "dx", atol=1e-9)  # B = 1: both gradients are exactly 0
    assert_close(dy, dy_o, 4e-3, "dy", atol=1e-9)


@pytest.mark.parametrize("mode,B", [("f32", 700), ("bf16", 4096), ("bf16", 8192 + 64)])
def test_clip_multi_tile_vs_oracle(ops, mode, B):
    from pclip import synth
    x, y = synth.clip_scaled_embeddings(B, 128, cfg_id=7, trained_like=True)
    loss_o, dx_o, dy_o, _, _ = R.clip_loss_and_grads(x.double(), y.double(), chunk=2048)
    loss, dx, dy, _, _ = ops.clip_loss_raw([x.cuda()], [y.cuda()], _mode(ops, mode), True)
    tol = TOL[mode]
    assert_close(loss[0], loss_o, tol, "loss")
    assert_close(dx, dx_o, tol * 3, "dx")
    assert_close(dy, dy_o, tol * 3, "dy")


@pytest.mark.parametrize("Ma,Nb,off", [(16, 16, 0), (129, 129, 0), (256, 256, 0), (257, 257, 0), (1000, 1000, 0),
                                       (4096, 4096, 0), (8256, 8256, 0), (512, 2048, 512), (300, 1200, 900),
                                       (2048, 16384, 6144)])
def test_single_pass_sumexp_vs_oracle(ops, Ma, Nb, off):
    """pclip_clip_sumexp (one sweep over S: row sums, column sums, diagonal; ref: training_utils.py:85-88) on square,
    ragged and row-sharded (Ma < Nb, diag_offset) shapes against float64 on the same bf16-rounded operands.  Row sums
    are fp32 sums of fp32 exps; column sums are fp32 sums of bf16-rounded exps (relative error <= 2^-9 per term)."""
    from pclip import synth
    x, y = synth.clip_scaled_embeddings(Nb, 128, cfg_id=Nb + 7, trained_like=(Nb % 3 == 0))
    xb, yb = x.bfloat16().cuda(), y.bfloat16().cuda()
    plan = ops.ClipPlan(Ma, Nb, 128, ops.BF16, "cuda", diag_offset=off)
    plan.xl.copy_(xb[off:off + Ma]); plan.yl.copy_(yb[off:off + Ma])
    if not plan.local_is_full:
        plan.xf.copy_(xb); plan.yf.copy_(yb)
    plan.sumexp()
    s = xb[off:off + Ma].double() @ yb.double().t()
    e = s.exp()
    idx = torch.arange(Ma, device="cuda")
    assert float(((plan.rowsum.double() - e.sum(1)) / e.sum(1)).abs().max()) < 2e-5
    assert float(((plan.colsum_f.double() - e.sum(0)) / e.sum(0)).abs().max()) < 2.0 ** -9
    assert float((plan.diag.double() - s[idx, idx + off]).abs().max()) < 1e-5


def _oracle_lse_full(xr, yr, chunk=2048):
    """Row and column log-sum-exp of S = xr yr^T at full size on the CPU: two chunked passes over the 65 536^2 logits
    (fp32 GEMM + exp per chunk, float64 accumulation of the sums), ref: modules/training_utils.py:85-88."""
    B = xr.shape[0]
    R_ = torch.empty(B, dtype=torch.float64)
    C_ = torch.zeros(B, dtype=torch.float64)
    yt = yr.t().contiguous()
    for i0 in range(0, B, chunk):
        e = (xr[i0:i0 + chunk] @ yt).exp_()
        R_[i0:i0 + chunk] = e.sum(1, dtype=torch.float64)
        C_ += e.sum(0, dtype=torch.float64)
    return R_.log(), C_.log()


def test_clip_full_size_properties(ops):
    """cfg 4 (B = 65 536, bf16; the gradient runs with 2 column splits + grad_reduce_kernel at this size).
    (a) all-zero embeddings: every logit is 0, so loss = ln B and both gradients vanish;
    (b) against an INDEPENDENT oracle: row and column LSEs of all 65 536 rows from two chunked CPU passes over the
        bf16-rounded inputs, the loss from them, and 256-row blocks of BOTH gradients from the closed form
        dS = (E/R + E/C - 2I) / 2B with the oracle's R and C (ref: modules/training_utils.py:85-89)."""
    B, D = 65536, 128
    z = torch.zeros(B, D, device="cuda")
    loss, dx, dy, _, _ = ops.clip_loss_raw([z], [z], ops.BF16, True)
    assert abs(float(loss[0]) - math.log(B)) < 1e-4
    assert float(dx.abs().max()) == 0.0 and float(dy.abs().max()) == 0.0
    from pclip import synth
    x, y = synth.clip_scaled_embeddings(B, D, cfg_id=4)
    loss, dx, dy, rl, cl = ops.clip_loss_raw([x.cuda()], [y.cuda()], ops.BF16, True, want_lse=True)
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    xr, yr = R.bf16_round(x), R.bf16_round(y)
    rl_o, cl_o = _oracle_lse_full(xr, yr)
    assert_close(rl.cpu(), rl_o, 1e-5, "row_lse (all rows)")
    assert_close(cl.cpu(), cl_o, 2e-5, "col_lse (all columns)")   # column sums add bf16-rounded exps (unbiased)
    diag = (xr.double() * yr.double()).sum(-1)
    loss_o = (rl_o + cl_o - 2 * diag).sum() / (2 * B)
    assert_close(loss[0], loss_o, 1e-5, "loss vs oracle")
    # gradient blocks: rows 256 k' of x (dX = G y) and of y (dY = G^T x), oracle normalisers
    idx = torch.arange(0, B, B // 256)
    x64, y64 = xr.double(), yr.double()
    s = x64[idx] @ y64.t()
    g = torch.exp(s - rl_o[idx, None]) + torch.exp(s - cl_o[None, :])
    dx_o = (g @ y64 - 2 * y64[idx]) / (2 * B)
    assert_close(dx.cpu()[idx], dx_o, 4e-3, "dx block")
    st = y64[idx] @ x64.t()                                           # S^T rows: [j, i]
    gt = torch.exp(st - rl_o[None, :]) + torch.exp(st - cl_o[idx, None])
    dy_o = (gt @ x64 - 2 * x64[idx]) / (2 * B)
    assert_close(dy.cpu()[idx], dy_o, 4e-3, "dy block")


def test_no_cpu_fallback(ops):
    from pclip import _lib
    with pytest.raises(_lib.PclipError):
        ops.clip_loss(torch.randn(4, 128), torch.randn(4, 128))


def test_gradcache_more_than_32_leaves(ops):
    """the pack kernel takes 32 chunks; train_gc(accumulated_batches=40) must still work like the
    reference's cat_input_tensor, which accepts any number of leaves."""
    from pclip import synth
    x, y = synth.clip_scaled_embeddings(40 * 8, 128, cfg_id=91)
    lo, dxo, dyo, _, _ = R.clip_loss_and_grads(x.double(), y.double())
    xs = [t.cuda().requires_grad_() for t in x.split(8)]
    ys = [t.cuda().requires_grad_() for t in y.split(8)]
    loss = ops.clip_loss(xs, ys, compute=ops.F32)
    loss.backward()
    assert_close(loss, lo, 1e-5, "loss, 40 leaves")
    assert_close(torch.cat([t.grad for t in xs]), dxo, 3e-5, "dx, 40 leaves")
    assert_close(torch.cat([t.grad for t in ys]), dyo, 3e-5, "dy, 40 leaves")


@pytest.mark.parametrize("name", ["clip_b16", "clip_b37", "clip_b256"])
def test_small_fused_kernel_against_reference_golden(ops, golden, name):
    """The one-launch small-batch kernel (pclip_clip_small_fused: fp32 arithmetic, cooperative grid barrier) against
    the reference's own loss and gradients, to the fp32 tolerance -- through autograd and through the launch counter."""
    g = golden(name)
    pep, rec = g["pep"].cuda().requires_grad_(), g["rec"].cuda().requires_grad_()
    ops.reset_launch_count()
    loss = ops.clip_loss(pep, rec.t().t(), compute=ops.BF16)     # B <= 512: the fused kernel serves both compute modes
    assert ops.launch_count() == 1
    loss.backward()
    assert_close(loss, g["loss"], 1e-5, "loss")
    assert_close(pep.grad, g["dpep"], 3e-5, "dpep")
    assert_close(rec.grad, g["drec"], 3e-5, "drec")


def test_gradcache_direct_path(ops, golden):
    """train_gc's loss step (ref: modules/training_utils.py:137-139) as ONE launch: 16 + 16 leaves, GradScaler's
    device-side scale 65536, gradients written straight into the leaves' .grad; against the reference's golden."""
    g = golden("clip_gradcache_16x16")
    xs = [t.cuda().requires_grad_() for t in g["x"]]
    ys = [t.cuda().requires_grad_() for t in g["y"]]
    scale = torch.tensor(65536.0, device="cuda")
    ops.reset_launch_count()
    loss = ops.clip_loss_gc_direct(xs, ys, scale)
    assert ops.launch_count() == 1 and loss.grad_fn is None
    assert_close(loss, g["loss"], 1e-5, "loss")
    assert_close(torch.stack([t.grad for t in xs]), g["dx"], 3e-5, "leaf grads x")
    assert_close(torch.stack([t.grad for t in ys]), g["dy"], 3e-5, "leaf grads y")
    # a second group accumulates into existing .grad like autograd would
    loss2 = ops.clip_loss_gc_direct(xs, ys, scale)
    assert_close(torch.stack([t.grad for t in xs]), 2 * g["dx"], 3e-5, "accumulated")
    # ragged leaves, bf16 leaves and a batch too large for the fused kernel take the same entry point
    x, y = torch.randn(600, 128).cuda(), torch.randn(600, 128).cuda()
    x, y = x / x.norm(dim=-1, keepdim=True), y / y.norm(dim=-1, keepdim=True)
    big = [t.requires_grad_() for t in (x[:100].clone(), x[100:].clone())], [y.clone().requires_grad_()]
    lb = ops.clip_loss_gc_direct(big[0], big[1])
    lo, dxo, dyo, _, _ = R.clip_loss_and_grads(R.bf16_round(x.cpu()).double(), R.bf16_round(y.cpu()).double())
    assert_close(lb, lo, 1e-2, "fallback loss")
    assert_close(torch.cat([t.grad for t in big[0]]), dxo, 1e-2, "fallback grads")
