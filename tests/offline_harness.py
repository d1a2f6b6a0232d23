"""Offline harness (TEST INFRASTRUCTURE): runs the reference's UNMODIFIED entry scripts -- main.py,
main_2protein.py, main_2protein_filip.py, topk.py, cosine_similarity.py -- against this repo's drop-in `modules`
package, with everything the scripts fetch from the network replaced by local stand-ins:

  * facebook/esm2_t30_150M_UR50D   -> a randomly initialised 2-layer EsmModel of the same width (640) and the 33-token ESM
                                     vocabulary (EsmModel.from_pretrained / EsmTokenizer.from_pretrained are patched)
  * Propedia / RCSB downloads, mmseqs2 -> seeded data/{peptide,receptor,protein1,protein2}.fasta + *DB_clustered.tsv written
                                     beforehand in the formats the loaders read (ref: modules/data_utils.py:99-106,143-150)
  * Bio, rcsbsearchapi, matplotlib -> stub modules (BioPython's SeqIO.parse is a ten-line FASTA reader)
  * the authors' checkpoints (hard-coded cluster paths, topk.py:42, cosine_similarity.py:41) -> load_state_dict of a
                                     missing file is a no-op

    python tests/offline_harness.py main_2protein.py [--pairs 400] [--backend cuda|oracle]

`--backend cuda` (default when a GPU is present) runs the real libpclip.so kernels; `--backend oracle` swaps pclip.ops for
the CPU oracle (tests/test_reference_scripts.py does that in the authoring container, which has the reference but no
GPU; the GPU box has a GPU but no reference checkout).  Nothing here is imported by the product.
"""
from __future__ import annotations

import importlib
import os
import runpy
import sys
import types

import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(REPO, "protein-clip_b200")
REFERENCE = os.environ.get("PCLIP_REFERENCE_ROOT", "/root/reference")
RESIDUES = "ACDEFGHIKLMNPQRSTVWY"
VOCAB = ["<cls>", "<pad>", "<eos>", "<unk>"] + list("LAGVSERTIDPKQNFYMHWCXBUZO.-") + ["<null_1>", "<mask>"]


# ---------------------------------------------------------------------------------------------------------
# data
# ---------------------------------------------------------------------------------------------------------
def make_data(workdir: str, pairs: int, seed: int = 0, len_a=(4, 14), len_b=(10, 40)):
    """Seeded FASTA + cluster files for both loaders; every sequence is unique and its own mmseqs cluster."""
    g = torch.Generator().manual_seed(seed)
    d = os.path.join(workdir, "data")
    os.makedirs(d, exist_ok=True)

    def seqs(n, lo, hi, tag):
        out = set()
        while len(out) < n:
            L = int(torch.randint(lo, hi + 1, (1,), generator=g))
            out.add("".join(RESIDUES[i] for i in torch.randint(0, 20, (L,), generator=g).tolist()))
        return sorted(out)

    a, b = seqs(pairs, *len_a, "a"), seqs(pairs, *len_b, "b")
    for stem_a, stem_b in (("peptide", "receptor"), ("protein1", "protein2")):
        with open(os.path.join(d, f"{stem_a}.fasta"), "w") as f:
            f.writelines(f">{stem_a}_{i}\n{s}\n" for i, s in enumerate(a))
        with open(os.path.join(d, f"{stem_b}.fasta"), "w") as f:
            f.writelines(f">{stem_b}_{i}\n{s}\n" for i, s in enumerate(b))
        with open(os.path.join(d, f"{stem_b}DB_clustered.tsv"), "w") as f:
            f.writelines(f"{stem_b}_{i}\t{stem_b}_{i}\n" for i in range(pairs))


# ---------------------------------------------------------------------------------------------------------
# stubs
# ---------------------------------------------------------------------------------------------------------
class _Anything(types.ModuleType):
    """A module whose every attribute is a callable that accepts anything and returns another such object."""

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        return _Callable(f"{self.__name__}.{name}")


class _Callable:
    def __init__(self, name):
        self._name = name

    def __call__(self, *a, **k):
        return _Callable(self._name + "()")

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        return _Callable(f"{self._name}.{name}")

    def __iter__(self):
        return iter(())


def _module(name, **attrs):
    m = sys.modules.get(name)
    if m is None:
        m = types.ModuleType(name)
        sys.modules[name] = m
    m.__dict__.update(attrs)
    return m


class _Record:
    def __init__(self, rid, seq):
        self.id, self.seq = rid, seq


def _fasta_parse(path, fmt="fasta"):
    rid, parts = None, []
    with open(path) as f:
        for line in f:
            line = line.strip()
            if line.startswith(">"):
                if rid is not None:
                    yield _Record(rid, "".join(parts))
                rid, parts = line[1:].split()[0], []
            elif line:
                parts.append(line)
    if rid is not None:
        yield _Record(rid, "".join(parts))


def install_stubs():
    try:
        from Bio import PDB, SeqIO  # noqa: F401
    except Exception:  # absent, or a partial stub another test left in sys.modules
        bio = _module("Bio")
        bio.SeqIO = _module("Bio.SeqIO", parse=_fasta_parse)
        pdb = _Anything("Bio.PDB")
        sys.modules["Bio.PDB"] = pdb
        bio.PDB = pdb
        _module("Bio.PDB.PDBExceptions", PDBConstructionWarning=type("PDBConstructionWarning", (Warning,), {}))
    for name in ("rcsbsearchapi", "rcsbsearchapi.search", "rcsbsearchapi.rcsb_attributes"):
        if name not in sys.modules:
            sys.modules[name] = _Anything(name)
    _module("rcsbsearchapi.const", CHEMICAL_ATTRIBUTE_SEARCH_SERVICE="text_chem", STRUCTURE_ATTRIBUTE_SEARCH_SERVICE="text")
    try:
        import matplotlib
        matplotlib.use("Agg")
        import matplotlib.pyplot  # noqa: F401
    except Exception:
        plt = _Anything("matplotlib.pyplot")

        def savefig(path, *a, **k):
            open(path, "wb").close()
        plt.savefig = savefig
        mpl = _module("matplotlib", use=lambda *a, **k: None)
        mpl.pyplot = plt
        sys.modules["matplotlib.pyplot"] = plt


class OfflineTokenizer:
    """The 33-token ESM vocabulary: <cls> + residues + <eos>, right-padded with <pad> (id 1), like EsmTokenizer."""
    tok = {t: i for i, t in enumerate(VOCAB)}

    @classmethod
    def from_pretrained(cls, name, *a, **k):
        return cls()

    def __call__(self, seqs, return_tensors="pt", padding=True):
        from transformers import BatchEncoding
        rows = [[0] + [self.tok.get(c, 3) for c in s] + [2] for s in seqs]
        T = max(len(r) for r in rows)
        ids = torch.tensor([r + [1] * (T - len(r)) for r in rows], dtype=torch.int64)
        return BatchEncoding({"input_ids": ids, "attention_mask": (ids != 1).to(torch.int64)})


def _offline_esm(name, *a, **k):
    from transformers import EsmConfig, EsmModel
    state = torch.get_rng_state()
    torch.manual_seed(1234)
    cfg = EsmConfig(vocab_size=33, hidden_size=640, num_hidden_layers=2, num_attention_heads=20, intermediate_size=1280,
                    position_embedding_type="rotary", token_dropout=True, pad_token_id=1, mask_token_id=32,
                    hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0)
    m = EsmModel(cfg, add_pooling_layer=False)
    torch.set_rng_state(state)
    return m


class _MissingCheckpoint(dict):
    pass


def patch_environment():
    """HF downloads, checkpoint paths of the authors' cluster and the GradScaler import the scripts use."""
    import transformers
    patches = []

    def setp(obj, name, value):
        patches.append((obj, name, getattr(obj, name)))
        setattr(obj, name, value)

    setp(transformers.EsmModel, "from_pretrained", staticmethod(_offline_esm))
    setp(transformers.EsmTokenizer, "from_pretrained", staticmethod(lambda *a, **k: OfflineTokenizer()))
    real_load = torch.load

    def load(path, *a, **k):
        if isinstance(path, (str, os.PathLike)) and not os.path.exists(path):
            print(f"[offline harness] checkpoint {path} is not available here: keeping the random initialisation")
            return _MissingCheckpoint()
        return real_load(path, *a, **k)

    setp(torch, "load", load)
    real_lsd = torch.nn.Module.load_state_dict

    def load_state_dict(self, sd, *a, **k):
        if isinstance(sd, _MissingCheckpoint):
            return None
        return real_lsd(self, sd, *a, **k)

    setp(torch.nn.Module, "load_state_dict", load_state_dict)

    def undo():
        for obj, name, old in reversed(patches):
            setattr(obj, name, old)
    return undo


def use_oracle_backend():
    """CPU stand-in for pclip.ops built on oracle/restate.py (differentiable torch): ONLY for running the scripts where
    there is no GPU.  Returns an undo function."""
    for p in (REPO, HOST, os.path.dirname(os.path.abspath(__file__))):
        if p not in sys.path:
            sys.path.insert(0, p)
    import _oracle_ops
    return _oracle_ops.install()


def run_script(script: str, workdir: str, backend: str | None = None, pairs: int = 400, seed: int = 0):
    """Run <reference>/<script> as __main__ with cwd = workdir.  Returns the exception the script ended with (None on a
    clean finish)."""
    path = os.path.join(REFERENCE, script)
    if not os.path.isfile(path):
        raise FileNotFoundError(f"{path}: the reference checkout is needed to run its scripts")
    backend = backend or ("cuda" if torch.cuda.is_available() else "oracle")
    make_data(workdir, pairs, seed)
    install_stubs()
    undo_env = patch_environment()
    undo_ops = use_oracle_backend() if backend == "oracle" else (lambda: None)
    old_path, old_cwd = list(sys.path), os.getcwd()
    dropped = {k: sys.modules.pop(k) for k in list(sys.modules) if k == "modules" or k.startswith("modules.")}
    sys.dont_write_bytecode = True
    # the drop-in first, the reference second: `modules` is a namespace package in both trees, so models /
    # training_utils* / visualizations / seed come from this repo and data_utils* from the reference
    sys.path[:0] = [HOST, REFERENCE]
    os.chdir(workdir)
    err = None
    try:
        runpy.run_path(path, run_name="__main__")
    except Exception as e:  # the caller decides which endings are acceptable (main.py:83 is a NameError upstream)
        import traceback
        err = e
        err.harness_traceback = traceback.format_exc()
    finally:
        os.chdir(old_cwd)
        sys.path[:] = old_path
        for k in [k for k in sys.modules if k == "modules" or k.startswith("modules.")]:
            sys.modules.pop(k)
        sys.modules.update(dropped)
        undo_ops()
        undo_env()
    return err


if __name__ == "__main__":
    import argparse
    import tempfile
    ap = argparse.ArgumentParser()
    ap.add_argument("script")
    ap.add_argument("--pairs", type=int, default=400)
    ap.add_argument("--backend", choices=["cuda", "oracle"], default=None)
    ap.add_argument("--workdir", default=None)
    a = ap.parse_args()
    wd = a.workdir or tempfile.mkdtemp(prefix="pclip_offline_")
    e = run_script(a.script, wd, a.backend, a.pairs)
    if e is not None:
        print(e.harness_traceback)
    print(f"[offline harness] {a.script} finished in {wd}: {'ok' if e is None else repr(e)}")
    sys.exit(0 if e is None else 1)
