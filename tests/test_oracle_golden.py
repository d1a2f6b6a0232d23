"""Pins oracle/restate.py against the golden vectors produced by the reference's own functions
(oracle/make_golden.py).  CPU only."""
import numpy as np
import pytest
import torch

from oracle import restate as R

F64 = torch.float64


def close(a, b, rtol=2e-5, atol=1e-6):
    a = torch.as_tensor(a, dtype=F64)
    b = torch.as_tensor(b, dtype=F64)
    scale = b.abs().max().clamp_min(1e-30)
    assert a.shape == b.shape
    err = (a - b).abs().max()
    assert err <= rtol * scale + atol * 0 + 1e-12, f"max err {err:.3e} vs scale {scale:.3e}"


def test_encoder_tail(golden):
    g = golden("encoder_tail", F64)
    close(R.normalize_scale(g["e"], g["tau"]), g["out"])
    de, dtau = R.normalize_scale_bwd(g["e"], g["tau"], g["g"])
    close(de, g["de"])
    close(dtau, g["dtau"])


@pytest.mark.parametrize("name", ["clip_b16", "clip_b37", "clip_b256"])
def test_clip_loss_and_grads(golden, name):
    g = golden(name, F64)
    pep = R.normalize_scale(g["e_pep"], g["tau"])
    rec = R.normalize_scale(g["e_rec"], g["tau"])
    close(pep, g["pep"])
    close(R.clip_loss(pep, rec.t()), g["loss"])
    # float32 literal arithmetic reproduces the reference's float32 result to a few ulp
    l32 = R.clip_loss(g["pep"].float(), g["rec"].float().t())
    close(l32, g["loss"], rtol=2e-6)
    loss, dx, dy, _, _ = R.clip_loss_and_grads(pep, rec, chunk=7)
    close(loss, g["loss"])
    close(dx, g["dpep"])
    close(dy, g["drec"])
    de_pep, dtau_p = R.normalize_scale_bwd(g["e_pep"], g["tau"], dx)
    de_rec, dtau_r = R.normalize_scale_bwd(g["e_rec"], g["tau"], dy)
    close(de_pep, g["de_pep"])
    close(de_rec, g["de_rec"])
    close(dtau_p + dtau_r, g["dtau"])


def test_clip_trained_like(golden):
    g = golden("clip_trained_like", F64)
    loss, dx, dy, _, _ = R.clip_loss_and_grads(g["pep"], g["rec"])
    close(loss, g["loss"])
    close(dx, g["dpep"])
    close(dy, g["drec"])


def test_gradcache_leaves(golden):
    g = golden("clip_gradcache_16x16", F64)
    xs, ys = list(g["x"]), list(g["y"])
    close(R.clip_loss_gc(xs, ys), g["loss"])
    loss, dx, dy, _, _ = R.clip_loss_and_grads(torch.cat(xs), torch.cat(ys), grad_output=float(g["grad_output"]))
    close(dx.view(16, 16, 128), g["dx"])
    close(dy.view(16, 16, 128), g["dy"])
    close(R.get_logits(xs, ys), g["exp_logits"])


def test_rank_of_positive(golden):
    g = golden("rank_of_positive")
    pep, rec = g["pep"], g["rec"]
    for k in range(pep.shape[0]):
        # the reference repeats query k against the FIRST 64 candidates and compares exp(logits)
        scores = (pep[k][None] @ rec.t()).exp()[0]
        assert R.rank_of_positive(scores, k) == int(g["ranks"][k])
    acc = R.topk_accuracy_from_ranks([int(r) for r in g["ranks"] if r < 64], 64)
    assert acc[-1] <= 1.0 and all(acc[i] <= acc[i + 1] for i in range(63))


def test_cosine(golden):
    g = golden("cosine_16", F64)
    close(R.cosine_matrix(g["pep"], g["rec"], g["tau"]), g["cos"])


@pytest.mark.parametrize("name", ["filip_b5_ragged", "filip_b4_full", "filip_b6_t70"])
def test_filip(golden, name):
    g = golden(name, F64)
    out = R.filip_loss_and_grads(g["hA"], g["hB"], g["maskA"], g["maskB"], g["tau"], normalize=True)
    close(out["simA"], g["simA"])
    close(out["simB"], g["simB"])
    close(out["loss"], g["loss"])
    close(out["dhA"], g["dhA"])
    close(out["dhB"], g["dhB"])
    close(out["dtau"], g["dtau"])
    close(R.filip_contrastive_loss(out["simA"], use_dcl=True), g["loss_dcl"])


def test_filip_forward_and_masked_mean(golden):
    g = golden("filip_forward_b3", F64)
    simA, simB = R.filip_similarity(R.filip_normalize(g["hA"]), R.filip_normalize(g["hB"]), g["maskA"], g["maskB"],
                                    g["tau"])
    close(simA, g["simA"])
    close(simB, g["simB"])
    m = golden("masked_mean")
    out = R.masked_mean(m["t"], m["mask"].bool(), dim=1)
    assert torch.equal(out, m["out"])


def test_topk_stable_ties():
    bank = torch.tensor([[1.0, 0.0], [0.0, 1.0], [1.0, 0.0], [0.5, 0.5], [1.0, 0.0]])
    q = torch.tensor([[1.0, 0.0], [0.0, 1.0]])
    idx, val = R.topk_stable(q, bank, 3)
    assert idx.tolist() == [[0, 2, 4], [1, 3, 0]]
    assert val[0].tolist() == [1.0, 1.0, 1.0]


def test_untrained_loss_is_ln_b():
    # soft anchor from the reference's notebooks/filip.ipynb cell 19: -log(1/16) = 2.7725887
    z = torch.zeros(16, 128, dtype=F64)
    assert abs(float(R.clip_loss(z, z.t())) - np.log(16)) < 1e-12


def test_encoder_head(golden):
    """The restated head (projection, per-residue FFN, masked mean, pooled FFN, normalise, temperature) against the
    reference's own Encoder / FILIPEncoder forward behind a fake ESM (eval mode), SURVEY.md 8(f)2."""
    g = golden("encoder_head", F64)
    sd = {k[3:]: v for k, v in g.items() if k.startswith("sd/")}
    close(R.encoder_head(g["hidden"], g["mask"], sd, g["tau"], pooled=True), g["pooled"])
    close(R.encoder_head(g["hidden"], g["mask"], sd, pooled=False), g["tokens"])
