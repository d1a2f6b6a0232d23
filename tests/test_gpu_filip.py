"""GPU parity of the FILIP token-wise late-interaction path (libpclip.so through the C ABI) against the golden vectors
generated from the reference's own ExtendedFILIP._filip_similarity_score + training_utils_filip._compute_loss, and
against the float64 oracle.  Tolerances are north_star's: 1e-5 relative in fp32, 1e-2 in bf16."""
import pytest
import torch

from oracle import restate as R

pytestmark = pytest.mark.gpu
TOL = {"f32": 1e-5, "bf16": 1e-2}


def rel_err(a, b):
    a = torch.as_tensor(a).detach().double().cpu()
    b = torch.as_tensor(b).detach().double().cpu()
    assert a.shape == b.shape, (a.shape, b.shape)
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


@pytest.fixture(scope="module")
def ops():
    from pclip import ops as o
    return o


def _mode(ops, name):
    return ops.F32 if name == "f32" else ops.BF16


def run_filip(ops, hA, hB, maskA, maskB, tau, mode, normalize):
    hA = hA.cuda().requires_grad_()
    hB = hB.cuda().requires_grad_()
    tau = tau.cuda().requires_grad_()
    nA = ops.normalize_scale(hA, None, eps=0.0) if normalize else hA
    nB = ops.normalize_scale(hB, None, eps=0.0) if normalize else hB
    simA, simB = ops.filip_similarity(nA, nB, maskA.cuda(), maskB.cuda(), tau, compute=mode)
    loss = ops.filip_loss(simA, simB)
    loss.backward()
    return simA, simB, loss, hA.grad, hB.grad, tau.grad


def test_filip_forward_golden(ops, golden):
    g = golden("filip_forward_b3")
    for name in ("f32", "bf16"):
        nA = ops.normalize_scale(g["hA"].cuda(), None, eps=0.0)     # ExtendedFILIP.forward, models.py:238-239
        nB = ops.normalize_scale(g["hB"].cuda(), None, eps=0.0)
        simA, simB = ops.filip_similarity(nA, nB, g["maskA"].cuda(), g["maskB"].cuda(), g["tau"].cuda(),
                                          compute=_mode(ops, name))
        assert rel_err(simA, g["simA"]) <= TOL[name]
        assert rel_err(simB, g["simB"]) <= TOL[name]


@pytest.mark.parametrize("name", ["filip_b4_full", "filip_b5_ragged", "filip_b6_t70"])
@pytest.mark.parametrize("mode", ["f32", "bf16"])
def test_filip_golden_fwd_bwd(ops, golden, name, mode):
    """ExtendedFILIP.forward's normalise (models.py:238-239) + _filip_similarity_score + _compute_loss, fwd + bwd."""
    g = golden(name)
    simA, simB, loss, dhA, dhB, dtau = run_filip(ops, g["hA"], g["hB"], g["maskA"], g["maskB"], g["tau"],
                                                 _mode(ops, mode), normalize=True)
    tol = TOL[mode]
    assert rel_err(simA, g["simA"]) <= tol
    assert rel_err(simB, g["simB"]) <= tol
    assert rel_err(loss, g["loss"]) <= tol
    assert rel_err(dtau, g["dtau"]) <= 3 * tol
    if mode == "f32":
        assert rel_err(dhA, g["dhA"]) <= 3 * tol
        assert rel_err(dhB, g["dhB"]) <= 3 * tol
    else:
        # bf16 rounding may flip near-tied maxima, which re-routes single gradient rows; compare in aggregate
        for got, want in ((dhA, g["dhA"]), (dhB, g["dhB"])):
            d = (got.double().cpu() - want.double())
            assert float(d.norm() / want.double().norm()) <= 5e-2


def _oracle_on_rounded(hA, hB, maskA, maskB, tau):
    nA, nB = R.filip_normalize(hA.double()), R.filip_normalize(hB.double())
    return R.filip_loss_and_grads(R.bf16_round(nA.float()).double(), R.bf16_round(nB.float()).double(), maskA, maskB,
                                  tau.double(), normalize=False)


@pytest.mark.parametrize("B,TA,TB,ragged", [(3, 130, 257, True), (6, 512, 300, True), (4, 256, 256, False)])
def test_filip_bf16_vs_oracle_on_rounded_inputs(ops, B, TA, TB, ragged):
    """Tensor-core path against the float64 oracle evaluated on the SAME bf16-rounded normalised tokens (isolates the
    kernel arithmetic from input rounding; the argmax then agrees except at fp32-accumulation near-ties)."""
    from pclip import synth
    hA, hB, mA, mB, tau = synth.filip_tokens(B, TA, TB, cfg_id=300 + TA, ragged=ragged)
    if ragged:
        mA[0, 1:] = 0       # a single-token sample
        mB[1, 5:] = 0
    o = _oracle_on_rounded(hA, hB, mA, mB, tau)
    nA = R.filip_normalize(hA.double()).float().cuda().requires_grad_()
    nB = R.filip_normalize(hB.double()).float().cuda().requires_grad_()
    t = tau.cuda().requires_grad_()
    simA, simB = ops.filip_similarity(nA, nB, mA.cuda(), mB.cuda(), t, compute=ops.BF16)
    assert rel_err(simA, o["simA"]) <= 1e-5
    assert rel_err(simB, o["simB"]) <= 1e-5
    loss = ops.filip_loss(simA, simB)
    assert rel_err(loss, o["loss"]) <= 1e-5
    loss.backward()
    assert rel_err(t.grad, o["dtau"]) <= 1e-4
    for got, want in ((nA.grad, o["dhA"]), (nB.grad, o["dhB"])):
        d = got.double().cpu() - want
        assert float(d.norm() / want.norm()) <= 1e-2      # bf16-rounded weights in the one-hot operand
    # pads receive exactly zero gradient
    for gr, m in ((nA.grad, mA), (nB.grad, mB)):
        pad = gr[~m.bool().cuda()]
        assert pad.numel() == 0 or float(pad.abs().max()) == 0.0


@pytest.mark.parametrize("mode", ["f32", "bf16"])
def test_filip_rectangular_row_shard(ops, mode):
    """Ba != Bb with a diagonal offset: what one rank of an A-row-sharded job computes."""
    from pclip import synth
    hA, hB, mA, mB, tau = synth.filip_tokens(6, 70, 90, cfg_id=77, ragged=False)
    nA, nB = R.filip_normalize(hA.double()).float(), R.filip_normalize(hB.double()).float()
    full = ops.filip_similarity(nA.cuda(), nB.cuda(), mA.cuda(), mB.cuda(), tau.cuda(), compute=_mode(ops, mode))
    part = ops.filip_similarity(nA[2:4].cuda(), nB.cuda(), mA[2:4].cuda(), mB.cuda(), tau.cuda(), compute=_mode(ops, mode))
    assert torch.equal(part[0], full[0][2:4]) and torch.equal(part[1], full[1][2:4])
    l_full = ops.filip_loss(*full)
    l_part = sum(float(ops.filip_loss(*ops.filip_similarity(nA[i:i + 2].cuda(), nB.cuda(), mA[i:i + 2].cuda(), mB.cuda(),
                                                            tau.cuda(), compute=_mode(ops, mode)), diag_offset=i)) * 2
                 for i in (0, 2, 4)) / 6
    assert abs(l_part - float(l_full)) <= 1e-5 * abs(float(l_full))
