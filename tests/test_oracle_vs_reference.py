"""The oracle against the LIVE reference (authoring container only: skipped where /root/reference is absent, e.g. on the
GPU box, where tests/test_oracle_golden.py checks the same functions against the committed golden vectors instead).
Fresh seeds on every parametrisation, so the restatement is exercised beyond the committed fixtures."""
import pytest
import torch

from oracle import ref_loader, restate as R

pytestmark = pytest.mark.skipif(not ref_loader.available(), reason="reference checkout not present")


@pytest.fixture(scope="module")
def ref():
    return ref_loader.load()


def rel(a, b):
    a, b = torch.as_tensor(a).detach().double(), torch.as_tensor(b).detach().double()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


@pytest.mark.parametrize("B,seed", [(7, 1), (16, 2), (53, 3)])
def test_clip_loss_and_gradients(ref, B, seed):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(B, 128, generator=g).requires_grad_()
    y = torch.randn(B, 128, generator=g).requires_grad_()
    tau = torch.tensor(0.7, requires_grad=True)
    import torch.nn.functional as F
    pep = F.normalize(x, dim=-1) * torch.exp(tau / 2)          # ref: modules/models.py:50-51
    rec = F.normalize(y, dim=-1) * torch.exp(tau / 2)
    loss = ref.training_utils._contrastive_loss(pep, rec.t())
    loss.backward()
    xo, yo, to = x.detach().double(), y.detach().double(), tau.detach().double()
    po, ro = R.normalize_scale(xo, to), R.normalize_scale(yo, to)
    lo, dpo, dro, _, _ = R.clip_loss_and_grads(po, ro)
    assert rel(lo, loss) <= 1e-6
    dx, dt1 = R.normalize_scale_bwd(xo, to, dpo)
    dy, dt2 = R.normalize_scale_bwd(yo, to, dro)
    assert rel(dx, x.grad) <= 1e-5 and rel(dy, y.grad) <= 1e-5 and rel(dt1 + dt2, tau.grad) <= 1e-5
    assert rel(R.clip_loss_gc(list(po.float().split(4)), list(ro.float().split(4))),
               ref.training_utils._contrastive_loss_gc(list(pep.detach().split(4)), list(rec.detach().split(4)))) <= 1e-6
    assert rel(R.get_logits(po.float(), ro.float()), ref.training_utils.get_logits(pep.detach(), rec.detach())) <= 1e-6


@pytest.mark.parametrize("B,TA,TB,seed", [(3, 9, 14, 4), (5, 33, 20, 5)])
def test_filip_similarity_loss_and_gradients(ref, B, TA, TB, seed):
    g = torch.Generator().manual_seed(seed)
    hA = torch.randn(B, TA, 128, generator=g).requires_grad_()
    hB = torch.randn(B, TB, 128, generator=g).requires_grad_()
    mA = (torch.rand(B, TA, generator=g) < 0.7).long()          # arbitrary masks, as the reference accepts
    mB = (torch.rand(B, TB, generator=g) < 0.7).long()
    mA[:, 0] = 1
    mB[:, 0] = 1
    tau = torch.tensor(1.3, requires_grad=True)
    M = ref.models.ExtendedFILIP
    nA = hA / hA.norm(dim=-1, keepdim=True)                     # ref: modules/models.py:238-239
    nB = hB / hB.norm(dim=-1, keepdim=True)
    sA, sB = M._filip_similarity_score(nA, nB, mA, mB, tau)
    loss = ref.training_utils_filip._compute_loss(sA, sB)
    loss.backward()
    o = R.filip_loss_and_grads(hA.detach().double(), hB.detach().double(), mA, mB, tau.detach().double())
    assert rel(o["simA"], sA) <= 1e-5 and rel(o["simB"], sB) <= 1e-5 and rel(o["loss"], loss) <= 1e-6
    assert rel(o["dhA"], hA.grad) <= 1e-4 and rel(o["dhB"], hB.grad) <= 1e-4 and rel(o["dtau"], tau.grad) <= 1e-4
    t = torch.randn(4, 6, generator=g)
    mk = torch.rand(4, 6, generator=g) < 0.5
    assert rel(R.masked_mean(t, mk), M._masked_mean(t, mk)) <= 1e-6
