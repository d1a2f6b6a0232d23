"""GPU parity of the fused encoder head (pclip_encoder_head_fwd through the C ABI) against the golden
vectors generated from the reference's own Encoder / FILIPEncoder forward, against the float64 oracle at the scripts'
width (ESM-2 150M: 640), and against the drop-in modules' stock nn.Sequential path.  fp32: 1e-5 relative."""
import types

import pytest
import torch

from oracle import restate as R

pytestmark = pytest.mark.gpu
TOL = 1e-5
DEV = "cuda"


def rel(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    assert a.shape == b.shape, (a.shape, b.shape)
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


class FakeEsm(torch.nn.Module):
    def __init__(self, hidden):
        super().__init__()
        self.hidden, self.calls = hidden, 0

    def forward(self, input_ids=None, attention_mask=None):
        self.calls += 1
        return types.SimpleNamespace(last_hidden_state=self.hidden)


def _encoder(kind, Din, h1, h2, hidden, seed):
    from modules import models
    torch.manual_seed(seed)
    enc = getattr(models, kind)(Din, 128, h1, h2, 0.1, FakeEsm(hidden)).to(DEV).eval()
    with torch.no_grad():
        for p in enc.parameters():
            if p.dim() == 1:
                p.add_(0.1 * torch.randn_like(p))
    return enc


def test_head_matches_reference_golden(golden):
    from modules import models
    g = golden("encoder_head")
    hidden, mask, tau = g["hidden"].to(DEV), g["mask"].to(DEV), g["tau"].to(DEV)
    sd = {k[3:]: v for k, v in g.items() if k.startswith("sd/")}
    for kind, want in (("Encoder", g["pooled"]), ("FILIPEncoder", g["tokens"])):
        enc = getattr(models, kind)(40, 128, 3, 2, 0.1, FakeEsm(hidden)).to(DEV).eval()
        enc.load_state_dict(sd, strict=False)
        seq = {"input_ids": torch.zeros_like(mask), "attention_mask": mask, "temperature": tau}
        with torch.no_grad():
            assert enc._fused_head_ok(enc.embedding_ffn if kind == "Encoder" else None)
            out = enc(seq)
        out = out[0] if kind == "FILIPEncoder" else out
        assert rel(out, want) <= TOL, kind
    assert torch.isfinite(g["pooled"][2]).all()      # the all-pad sample: count clamped to 1, mean of nothing = 0


@pytest.mark.parametrize("Bs,T,Din,h1,h2", [(7, 77, 640, 2, 2), (3, 32, 640, 1, 1), (2, 513, 64, 4, 3), (16, 9, 24, 2, 2)])
def test_head_matches_oracle_and_stock_path(Bs, T, Din, h1, h2, monkeypatch):
    from pclip import ops
    g = torch.Generator().manual_seed(Bs * 1000 + T)
    hidden = torch.randn(Bs, T, Din, generator=g).to(DEV)
    mask = (torch.rand(Bs, T, generator=g) < 0.7).long()      # arbitrary (non-prefix) masks
    mask[0] = 1
    if Bs > 2:
        mask[1] = 0                                             # all-pad sample
    mask = mask.to(DEV)
    tau = torch.tensor(1.3, device=DEV)
    for kind in ("Encoder", "FILIPEncoder"):
        enc = _encoder(kind, Din, h1, h2, hidden, seed=Bs + T)
        seq = lambda: {"input_ids": torch.zeros_like(mask), "attention_mask": mask, "temperature": tau}  # noqa: E731
        with torch.no_grad():
            ops.reset_launch_count()
            fused = enc(seq())
            launches = ops.launch_count()
            monkeypatch.setenv("PCLIP_FUSED_HEAD", "0")
            stock = enc(seq())
            monkeypatch.delenv("PCLIP_FUSED_HEAD")
        if kind == "FILIPEncoder":
            fused, stock = fused[0], stock[0]
        assert launches == (2 if kind == "Encoder" else 1), (kind, launches)
        sd = {k: v.double().cpu() for k, v in enc.state_dict().items()}
        want = R.encoder_head(hidden.double().cpu(), mask.cpu(), sd, tau.double().cpu(), pooled=kind == "Encoder")
        assert rel(fused, want) <= TOL, (kind, rel(fused, want))
        assert rel(stock, want) <= TOL, (kind, rel(stock, want))    # the stock path is within the same bound
        if kind == "FILIPEncoder":                                   # every residue row, pads included, as the reference
            assert float((fused.double().cpu() - want).abs().max()) <= TOL * float(want.abs().max())


def test_training_and_grad_paths_keep_the_stock_modules():
    hidden = torch.randn(2, 8, 40, device=DEV)
    enc = _encoder("Encoder", 40, 2, 2, hidden, seed=1)
    assert not enc._fused_head_ok(enc.embedding_ffn)                 # grad enabled
    enc.train()
    with torch.no_grad():
        assert not enc._fused_head_ok(enc.embedding_ffn)             # training mode (GradCache's first pass)
    enc.eval()
    with torch.no_grad():
        assert enc._fused_head_ok(enc.embedding_ffn)


def test_head_rejects_unsupported_shapes():
    from pclip import _lib as L, ops
    lin = torch.nn.Linear(42, 128).to(DEV)
    ffn = torch.nn.Sequential(torch.nn.Linear(128, 128)).to(DEV)
    assert not ops.encoder_head_supported(torch.empty(0, 42, device=DEV), lin, ffn, None)     # Din % 4 != 0
    with pytest.raises(L.PclipError, match="multiple of 4"):
        ops.encoder_head(torch.randn(1, 4, 42, device=DEV), torch.ones(1, 4, dtype=torch.int64, device=DEV), lin, ffn)
