"""Import the reference's own Python modules from /root/reference (authoring container only; TEST INFRASTRUCTURE).

The reference checkout does not exist on the GPU box, so nothing marked `gpu`, smoke() or bench.py uses this
file; it is used by oracle/make_golden.py (to write tests/golden/*.npz) and by tests/test_oracle_vs_reference.py
(skipped when the checkout is absent).

Shims needed to import (nothing in the reference is modified):
  * grad_cache            -> protein-clip_b200/grad_cache (upstream submodule is empty)
  * matplotlib / Bio      -> empty stub modules (only modules/visualizations.py imports them, for plotting)
"""
from __future__ import annotations

import importlib
import os
import sys
import types
import warnings

REFERENCE_ROOT = os.environ.get("PCLIP_REFERENCE_ROOT", "/root/reference")
_REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_HOST = os.path.join(_REPO, "protein-clip_b200")


def available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "modules", "training_utils.py"))


def _stub(name: str, **attrs):
    if name in sys.modules:
        return sys.modules[name]
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


def load():
    """Returns a namespace with the reference's models / training_utils / training_utils_filip /
    visualizations modules, imported under the package name `refmodules` so they never collide with the
    product's drop-in `modules` package."""
    if not available():
        raise RuntimeError(f"reference checkout not found at {REFERENCE_ROOT}")
    sys.dont_write_bytecode = True  # the mount is read-only
    if _HOST not in sys.path:
        sys.path.insert(0, _HOST)  # grad_cache shim
    try:
        import matplotlib.pyplot  # noqa: F401
    except Exception:
        mpl = _stub("matplotlib")
        mpl.pyplot = _stub("matplotlib.pyplot")
    try:
        import Bio  # noqa: F401
    except Exception:
        bio = _stub("Bio")
        bio.SeqIO = _stub("Bio.SeqIO")
    pkg = types.ModuleType("refmodules")
    pkg.__path__ = [os.path.join(REFERENCE_ROOT, "modules")]
    sys.modules["refmodules"] = pkg
    ns = types.SimpleNamespace()
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        for name in ("models", "training_utils", "training_utils_filip", "visualizations"):
            setattr(ns, name, importlib.import_module(f"refmodules.{name}"))
    return ns
