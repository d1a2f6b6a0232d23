"""Generate tests/golden/*.npz by running the REFERENCE'S OWN functions (imported from /root/reference) on
seeded synthetic inputs.  Run in the authoring container:  python -m oracle.make_golden

Each .npz stores the inputs next to the reference outputs, so the fixtures do not depend on RNG stability.
TEST INFRASTRUCTURE (see oracle/__init__.py).
"""
from __future__ import annotations

import os
import sys
import types

import numpy as np
import torch

_REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, _REPO)
sys.path.insert(0, os.path.join(_REPO, "protein-clip_b200"))

from oracle import ref_loader  # noqa: E402
from pclip import synth  # noqa: E402

OUT = os.path.join(_REPO, "tests", "golden")


def _np(d):
    return {k: (v.detach().cpu().numpy() if isinstance(v, torch.Tensor) else np.asarray(v)) for k, v in d.items()}


def _save(name, **arrs):
    os.makedirs(OUT, exist_ok=True)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **_np(arrs))
    print("wrote", name, {k: tuple(np.asarray(v).shape) for k, v in _np(arrs).items()})


class _FakeEsm(torch.nn.Module):
    """Stands in for the frozen HF EsmModel: returns a fixed [B, T, input_dim] hidden state."""

    def __init__(self, hidden):
        super().__init__()
        self.hidden = hidden

    def forward(self, input_ids=None, attention_mask=None):
        return types.SimpleNamespace(last_hidden_state=self.hidden)


def encoder_tail(ref):
    """models.Encoder.forward end to end on a fake ESM; the golden pair is (embedding_ffn output -> result),
    i.e. lines models.py:50-51, plus autograd gradients through them."""
    torch.manual_seed(7)
    B, T, Din, D = 16, 9, 24, 128
    hidden = torch.randn(B, T, Din)
    enc = ref.models.Encoder(Din, D, 2, 2, 0.0, _FakeEsm(hidden)).eval()
    grabbed = {}
    enc.embedding_ffn.register_forward_hook(lambda m, i, o: grabbed.__setitem__("e", o))
    tau = torch.tensor(0.7, requires_grad=True)
    mask = synth.prefix_mask(torch.randint(2, T + 1, (B,)), T)
    out = enc({"input_ids": torch.zeros(B, T, dtype=torch.long), "attention_mask": mask, "temperature": tau})
    e = grabbed["e"]
    g = torch.randn(B, D)
    de, dtau = torch.autograd.grad(out, [e, tau], g)
    _save("encoder_tail", e=e, tau=tau, out=out, g=g, de=de, dtau=dtau)


def clip_cases(ref):
    tu = ref.training_utils
    for name, B, cfg in (("clip_b16", 16, 1), ("clip_b37", 37, 11), ("clip_b256", 256, 12)):
        e_pep, e_rec, tau = synth.clip_raw_embeddings(B, 128, cfg)
        e_pep.requires_grad_()
        e_rec.requires_grad_()
        tau = tau.clone().requires_grad_()
        pep = torch.nn.functional.normalize(e_pep, dim=-1) * torch.exp(tau / 2)   # what Encoder.forward returns
        rec = torch.nn.functional.normalize(e_rec, dim=-1) * torch.exp(tau / 2)
        pep.retain_grad()
        rec.retain_grad()
        loss = tu._contrastive_loss(pep, rec.t())
        loss.backward()
        _save(name, e_pep=e_pep, e_rec=e_rec, tau=tau, pep=pep, rec=rec, loss=loss, dpep=pep.grad, drec=rec.grad,
              de_pep=e_pep.grad, de_rec=e_rec.grad, dtau=tau.grad)
    # trained-like embeddings (diagonal dominates), larger tau
    x, y = synth.clip_scaled_embeddings(64, 128, 13, tau=2.0, trained_like=True)
    x.requires_grad_()
    y.requires_grad_()
    loss = tu._contrastive_loss(x, y.t())
    loss.backward()
    _save("clip_trained_like", pep=x, rec=y, loss=loss, dpep=x.grad, drec=y.grad)


def gradcache_case(ref):
    tu = ref.training_utils
    xs, ys = synth.gradcache_leaves(16, 16, 128, 2)
    xs = [t.clone().requires_grad_() for t in xs]
    ys = [t.clone().requires_grad_() for t in ys]
    loss = tu._contrastive_loss_gc(xs, ys)
    (loss * 65536.0).backward()                      # GradScaler.scale(loss).backward(), training_utils.py:139
    logits = tu.get_logits([t.detach() for t in xs], [t.detach() for t in ys])
    _save("clip_gradcache_16x16", x=torch.stack([t.detach() for t in xs]), y=torch.stack([t.detach() for t in ys]),
          loss=loss, dx=torch.stack([t.grad for t in xs]), dy=torch.stack([t.grad for t in ys]),
          exp_logits=logits, grad_output=torch.tensor(65536.0))


class _Tok(dict):
    def to(self, device):
        return self


def rank_case(ref):
    """eval_gc_allrec_onepep end to end with stub encoders: 'sequences' are row indices into fixed embedding
    tables, the tokenizer passes them through."""
    tu = ref.training_utils
    g = torch.Generator().manual_seed(99)
    n = 64
    pep_table = synth.unit_rows(n, 128, g, scale=float(np.exp(0.5)))
    rec_table = pep_table + 0.8 * torch.randn(n, 128, generator=g) / np.sqrt(128)
    rec_table = rec_table / rec_table.norm(dim=-1, keepdim=True) * float(np.exp(0.5))

    def tokenizer(seqs, return_tensors="pt", padding=True):
        return _Tok(input_ids=torch.tensor([int(s) for s in seqs]))

    class Enc(torch.nn.Module):
        def __init__(self, table):
            super().__init__()
            self.table = table

        def forward(self, seq):
            return self.table[seq["input_ids"]]

    model = types.SimpleNamespace(pep_encoder=Enc(pep_table), rec_encoder=Enc(rec_table),
                                  temperature=torch.tensor(1.0), eval=lambda: None)
    loader = [(tuple(str(i) for i in range(b, b + 16)), tuple(str(i) for i in range(b, b + 16)))
              for b in range(0, n, 16)]
    ranks = [tu.eval_gc_allrec_onepep(model, loader, "cpu", tokenizer, agg_batches=4, k=k)[0] for k in range(n)]
    _save("rank_of_positive", pep=pep_table, rec=rec_table, ranks=np.asarray(ranks, dtype=np.int64))


def cosine_case(ref):
    vis = ref.visualizations
    pep, rec = synth.clip_scaled_embeddings(16, 128, 21, tau=1.3)
    model = lambda p, r: (pep, rec)  # noqa: E731
    model.temperature = torch.tensor(1.3)
    cos = vis._compute_embedding_cosine_similarities(model, None, None)
    _save("cosine_16", pep=pep, rec=rec, tau=model.temperature, cos=cos)


def filip_cases(ref):
    tf = ref.training_utils_filip
    FIL = ref.models.ExtendedFILIP
    for name, B, TA, TB, cfg, ragged in (("filip_b5_ragged", 5, 12, 20, 31, True), ("filip_b4_full", 4, 16, 16, 32, False),
                                         ("filip_b6_t70", 6, 70, 45, 33, True)):
        g = torch.Generator().manual_seed(1000 + cfg)
        hA = torch.randn(B, TA, 128, generator=g, requires_grad=True)
        hB = torch.randn(B, TB, 128, generator=g, requires_grad=True)
        if ragged:
            lenA = torch.randint(1, TA + 1, (B,), generator=g)
            lenB = torch.randint(1, TB + 1, (B,), generator=g)
            lenA[0] = TA
            lenB[-1] = TB
            lenA[-1] = 1
        else:
            lenA = torch.full((B,), TA)
            lenB = torch.full((B,), TB)
        mA, mB = synth.prefix_mask(lenA, TA), synth.prefix_mask(lenB, TB)
        tau = torch.tensor(0.9, requires_grad=True)
        nA = hA / hA.norm(dim=-1, keepdim=True)                    # models.py:238-239
        nB = hB / hB.norm(dim=-1, keepdim=True)
        simA, simB = FIL._filip_similarity_score(nA, nB, mA, mB, tau)
        loss = tf._compute_loss(simA, simB)
        dhA, dhB, dtau = torch.autograd.grad(loss, [hA, hB, tau])
        _save(name, hA=hA, hB=hB, maskA=mA, maskB=mB, tau=tau, simA=simA, simB=simB, loss=loss, dhA=dhA, dhB=dhB,
              dtau=dtau, loss_dcl=tf._contrastive_loss(simA.detach(), use_dcl=True))
    # the full ExtendedFILIP.forward with stubbed encoders (checks normalise + score wiring + dict mutation)
    model = FIL(8, 128, 2, 2, 0.0, _FakeEsm(torch.zeros(1)))
    g = torch.Generator().manual_seed(77)
    hA, hB = torch.randn(3, 7, 128, generator=g), torch.randn(3, 9, 128, generator=g)
    mA, mB = synth.prefix_mask(torch.tensor([7, 3, 5]), 7), synth.prefix_mask(torch.tensor([2, 9, 4]), 9)

    class Fixed(torch.nn.Module):
        def __init__(self, h, m):
            super().__init__()
            self.h, self.m = h, m

        def forward(self, seq):
            return self.h, self.m

    model.pep_encoder = Fixed(hA, mA)
    model.rec_encoder = Fixed(hB, mB)
    simA, simB = model({}, {})
    _save("filip_forward_b3", hA=hA, hB=hB, maskA=mA, maskB=mB, tau=model.temperature, simA=simA, simB=simB)
    # _masked_mean on its own
    t = torch.randn(4, 6, generator=g)
    m = synth.prefix_mask(torch.tensor([6, 0, 2, 1]), 6).bool()
    _save("masked_mean", t=t, mask=m, out=FIL._masked_mean(t, m, dim=1))


def encoder_head_case(ref):
    """models.Encoder.forward and models.FILIPEncoder.forward end to end behind a fake ESM (eval mode): the whole head,
    SURVEY.md 8(f)2.  Depths 3 / 2 so that more than one LayerNorm block is exercised; one all-pad sample (count
    clamp), one non-prefix mask."""
    torch.manual_seed(11)
    B, T, Din, D = 5, 45, 40, 128
    hidden = torch.randn(B, T, Din)
    mask = synth.prefix_mask(torch.tensor([45, 33, 0, 7, 20]), T)
    mask[4, 3] = 0
    mask[4, 30] = 1
    tau = torch.tensor(0.8)
    enc = ref.models.Encoder(Din, D, 3, 2, 0.1, _FakeEsm(hidden)).eval()
    for p in enc.parameters():          # LayerNorm weights / biases away from their 1 / 0 initialisation
        if p.dim() == 1:
            p.data.add_(0.1 * torch.randn_like(p))
    fil = ref.models.FILIPEncoder(Din, D, 3, 2, 0.1, _FakeEsm(hidden)).eval()
    fil.load_state_dict(enc.state_dict())
    ids = torch.zeros(B, T, dtype=torch.long)
    with torch.no_grad():
        pooled = enc({"input_ids": ids, "attention_mask": mask, "temperature": tau})
        tokens, m2 = fil({"input_ids": ids, "attention_mask": mask, "temperature": tau})
    assert torch.equal(m2, mask)
    sd = {"sd/" + k: v for k, v in enc.state_dict().items() if not k.startswith("esm_model")}
    _save("encoder_head", hidden=hidden, mask=mask, tau=tau, pooled=pooled, tokens=tokens, **sd)


def main():
    ref = ref_loader.load()
    torch.set_num_threads(1)  # fixed summation order for the fixtures
    encoder_tail(ref)
    clip_cases(ref)
    gradcache_case(ref)
    rank_case(ref)
    cosine_case(ref)
    filip_cases(ref)
    encoder_head_case(ref)


if __name__ == "__main__":
    main()
