"""CPU checks of the drop-in boundary: the C-ABI library exports every symbol include/pclip.h declares, the ctypes table
mirrors it, the `modules` package keeps the reference's signatures and state_dict keys (fixture generated from the
reference by oracle/make_interface_fixture.py), the grad_cache shim behaves like upstream, and nothing computes on CPU."""
import inspect
import json
import os
import re

import pytest
import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def iface():
    with open(os.path.join(REPO, "tests", "golden", "interface.json")) as f:
        return json.load(f)


def _params(sig_or_str):
    """[(name, default-or-None)] with annotations dropped; accepts an inspect.Signature or the fixture's string."""
    if isinstance(sig_or_str, str):
        body = sig_or_str.strip()[1:-1]
        out = []
        for part in re.split(r",\s*(?![^\[\]]*\])", body):
            part = part.strip()
            if not part:
                continue
            name = re.split(r"[:=]", part)[0].strip()
            default = part.split("=")[-1].strip() if "=" in part else None
            out.append((name, default))
        return out
    return [(p.name, None if p.default is inspect._empty else repr(p.default)) for p in sig_or_str.parameters.values()]


def _same(ours, ref):
    a, b = _params(inspect.signature(ours)), _params(ref)
    assert [n for n, _ in a] == [n for n, _ in b], (a, b)
    for (_, da), (_, db) in zip(a, b):
        assert (da is None) == (db is None), (a, b)
        if da is not None:
            assert float(da) == float(db) if re.fullmatch(r"[-0-9.e]+", db) else da == db, (a, b)


def test_header_symbols_are_exported_and_bound():
    from pclip import _lib
    hdr = open(os.path.join(REPO, "include", "pclip.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(pclip_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 24
    h = _lib.lib()
    for name in declared:
        assert hasattr(h, name), f"{name} declared in pclip.h but not exported"
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)
    assert h.pclip_version() >= 100


def test_no_cpu_path():
    from pclip import _lib, ops
    with pytest.raises(_lib.PclipError):
        ops.clip_loss(torch.randn(4, 128), torch.randn(4, 128))
    with pytest.raises(_lib.PclipError):
        ops.topk(torch.randn(4, 128), torch.randn(9, 128), 2)
    with pytest.raises(_lib.PclipError):
        ops.filip_similarity(torch.randn(2, 3, 128), torch.randn(2, 3, 128), torch.ones(2, 3), torch.ones(2, 3), 1.0)
    if not torch.cuda.is_available():   # the C entry points themselves refuse to run without a device
        h = _lib.lib()
        rc = h.pclip_scores(None, None, 1, 1, 128, 0, None, None, None)
        assert rc == -3 and b"no CUDA device" in h.pclip_last_error()


def test_module_signatures_match_reference(iface):
    from modules import models, training_utils, training_utils_filip, visualizations
    for name, ref in iface["training_utils"].items():
        _same(getattr(training_utils, name), ref)
    for name, ref in iface["training_utils_filip"].items():
        _same(getattr(training_utils_filip, name), ref)
    for name, ref in iface["visualizations"].items():
        _same(getattr(visualizations, name), ref)
    for name, ref in iface["models"].items():
        cls, attr = name.split(".")
        _same(getattr(getattr(models, cls), attr), ref)
    for fn in ("get_logits", "_call_model_gc", "_contrastive_loss_gc", "cached", "cat_input_tensor"):
        assert callable(getattr(training_utils, fn))


def test_every_attribute_the_scripts_use_exists(iface):
    """main*.py / topk.py / cosine_similarity.py must not meet an AttributeError on the modules this repo replaces."""
    import importlib
    for mod, attrs in iface["script_attrs"].items():
        m = importlib.import_module(f"modules.{mod}")
        for a in attrs:
            assert hasattr(m, a), f"modules.{mod}.{a} is used by the reference's scripts but missing here"


def test_cpu_plot_helpers(tmp_path):
    """The plotting helpers the scripts call at start-up and after the last epoch (CPU only; reference naming
    'figures/figure N.png' with one shared counter, ref: modules/visualizations.py:191-210)."""
    from modules import visualizations as vis
    (tmp_path / "protein1.fasta").write_text(">a\nACDE\nFG\n>b\nAAA\n")
    (tmp_path / "protein2.fasta").write_text(">c\nACDEFGHIK\n")
    (tmp_path / "protein2DB_clustered.tsv").write_text("c1\tc1\nc1\tx\nc2\tc2\n")
    assert vis._fasta_lengths(tmp_path / "protein1.fasta") == [6, 3]
    first = vis._save_plot(str(tmp_path))
    n0 = int(os.path.basename(first)[len("figure "):-4])
    vis.plot_protein_lengths(str(tmp_path), tmp_path)
    vis.plot_clustering(str(tmp_path), str(tmp_path), prefix="protein2")
    vis.plot_loss_curves(str(tmp_path), [3.0, 2.5], [3.1, 2.9], 256, 16)
    nxt = vis._save_plot(str(tmp_path))
    assert nxt.endswith(f"figure {n0 + 4}.png") and os.path.dirname(nxt) == str(tmp_path / "figures")
    made = sorted(os.listdir(tmp_path / "figures"))
    assert len(made) >= 3 and all(f.startswith("figure ") for f in made)


def test_state_dict_keys_match_reference(iface):
    from modules import models

    class FakeEsm(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.w = torch.nn.Parameter(torch.zeros(1))

    for cls in ("ExtendedCLIP", "ExtendedFILIP"):
        m = getattr(models, cls)(40, 128, 2, 2, 0.1, FakeEsm())
        keys = sorted(k for k in m.state_dict() if "esm_model" not in k)
        assert keys == iface["state_dict_keys"][cls]
        assert m.temperature.item() == 1.0 and isinstance(m.temperature, torch.nn.Parameter)


def test_grad_cache_shim_equals_direct_backward():
    """cached(): loss on detached leaves + closure replay == plain end-to-end backward, dropout masks included."""
    from grad_cache.functional import cached, cat_input_tensor
    torch.manual_seed(0)
    net = torch.nn.Sequential(torch.nn.Linear(8, 8), torch.nn.Dropout(0.5), torch.nn.Linear(8, 4))
    xs = [torch.randn(5, 8) for _ in range(3)]

    @cached
    def call(model, inp):
        return model(inp)

    @cat_input_tensor
    def loss_fn(reps):
        return (reps ** 2).sum()

    torch.manual_seed(123)
    pairs = [call(net, x) for x in xs]
    leaves = [p[0] for p in pairs]
    assert all(l.requires_grad and l.grad_fn is None for l in leaves)
    loss = loss_fn(leaves)
    loss.backward()
    for (leaf, closure) in pairs:
        closure(leaf)
    got = [p.grad.clone() for p in net.parameters()]
    net.zero_grad()
    torch.manual_seed(123)
    direct = sum((net(x) ** 2).sum() for x in xs)
    direct.backward()
    assert abs(float(direct) - float(loss)) < 1e-5
    for g, p in zip(got, net.parameters()):
        assert torch.allclose(g, p.grad, atol=1e-6)
