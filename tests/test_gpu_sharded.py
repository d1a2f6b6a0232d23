"""The row-sharded CLIP stages on ONE GPU: every rank's rectangular problem (local rows x all gathered rows, diagonal
offset) is run in turn with the gathers done by hand, and the union of the results is compared with the single-GPU
composite and the oracle.  (The NCCL path itself is exercised by `bench.py --gpus N`.)"""
import pytest
import torch

from oracle import restate as R

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("mode", ["f32", "bf16"])
@pytest.mark.parametrize("P,B", [(2, 512), (4, 1024), (8, 2048), (4, 4096)])
def test_emulated_ranks_match_single_gpu(P, B, mode, monkeypatch):
    from pclip import ops, synth
    monkeypatch.setenv("PCLIP_SMALL_FUSED", "0")   # the single-GPU reference must run the SAME (large-batch) kernels
    compute = ops.F32 if mode == "f32" else ops.BF16
    x, y = synth.clip_scaled_embeddings(B, 128, cfg_id=40 + P, trained_like=True)
    xd, yd = x.cuda(), y.cuda()
    Ma = B // P
    plans = [ops.ClipPlan(Ma, B, 128, compute, "cuda", diag_offset=r * Ma) for r in range(P)]
    for r, p in enumerate(plans):
        p.pack([xd[r * Ma:(r + 1) * Ma]], [yd[r * Ma:(r + 1) * Ma]])
    xf, yf = torch.cat([p.xl for p in plans]), torch.cat([p.yl for p in plans])      # the all-gather
    for p in plans:
        p.xf.copy_(xf)
        p.yf.copy_(yf)
        p.transpose_full()
    if plans[0].single_pass:       # pclip/sharded.py: one sweep per rank, column sums all-reduced over ranks
        for p in plans:
            p.sumexp()
        colsum = sum(p.colsum_f.clone() for p in plans)                              # the all-reduce of B floats
        for p in plans:
            p.colsum_f.copy_(colsum)
            p.take_local_colsum()
    else:
        for p in plans:
            p.rowsumexp()
    for p in plans:
        p.finalize(B)
    loss = sum(p.loss.clone() for p in plans)                                        # the all-reduce
    rinv, cinv = torch.cat([p.rinv for p in plans]), torch.cat([p.cinv for p in plans])
    for p in plans:
        p.rinv_f.copy_(rinv)
        p.cinv_f.copy_(cinv)
        p.grad(B)
    dx, dy = torch.cat([p.dx for p in plans]), torch.cat([p.dy for p in plans])
    l1, dx1, dy1, _, _ = ops.clip_loss_raw([xd], [yd], compute, True)
    tol = 1e-5 if mode == "f32" else 1e-2
    xo, yo = (x, y) if mode == "f32" else (R.bf16_round(x), R.bf16_round(y))
    lo, dxo, dyo, _, _ = R.clip_loss_and_grads(xo.double(), yo.double())
    assert abs(float(loss) - float(lo)) <= tol * abs(float(lo))
    assert abs(float(loss) - float(l1)) <= 1e-6 * abs(float(l1))
    for got, one, want in ((dx, dx1, dxo), (dy, dy1, dyo)):
        assert float((got.double().cpu() - want).abs().max() / want.abs().max()) <= (3e-5 if mode == "f32" else 4e-3)
        assert float((got - one).abs().max() / one.abs().max()) <= 1e-5


@pytest.mark.parametrize("P,B", [(2, 2048), (8, 4096)])
def test_emulated_fused_exchange_matches_single_gpu(P, B):
    """The one-collective exchange of pclip/sharded.py (record [C partial | R | S_ii] per rank, finalize_full on every
    rank) with the all-gather done by hand."""
    from pclip import ops, synth
    x, y = synth.clip_scaled_embeddings(B, 128, cfg_id=60 + P, trained_like=True)
    xd, yd = x.cuda(), y.cuda()
    Ma = B // P
    plans = [ops.ClipPlan(Ma, B, 128, ops.BF16, "cuda", diag_offset=r * Ma) for r in range(P)]
    assert all(p.single_pass and hasattr(p, "stats_l") for p in plans)
    for r, p in enumerate(plans):
        p.pack([xd[r * Ma:(r + 1) * Ma]], [yd[r * Ma:(r + 1) * Ma]])
    xf, yf = torch.cat([p.xl for p in plans]), torch.cat([p.yl for p in plans])
    for p in plans:
        p.xf.copy_(xf)
        p.yf.copy_(yf)
        p.sumexp()
    stats_all = torch.stack([p.stats_l for p in plans])                               # the single all-gather
    for p in plans:
        p.finalize_full(stats_all, B)
        p.grad(B)
    l1, dx1, dy1, _, _ = ops.clip_loss_raw([xd], [yd], ops.BF16, True)
    for p in plans:                                                                   # the loss is global on every rank
        assert abs(float(p.loss) - float(l1)) <= 1e-6 * abs(float(l1))
    dx, dy = torch.cat([p.dx for p in plans]), torch.cat([p.dy for p in plans])
    for got, one in ((dx, dx1), (dy, dy1)):
        assert float((got - one).abs().max() / one.abs().max()) <= 1e-5
