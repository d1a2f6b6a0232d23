"""`grad_cache.functional.cached` / `cat_input_tensor` with the upstream (luyug/GradCache) semantics.

Restated from the published algorithm (Gao et al. 2021, "Scaling Deep Contrastive Learning Batch Size under
Memory Limited Setup"); upstream source is not available offline, so this boundary is "parity unpinned" against
upstream -- tests/test_interface.py (test_grad_cache_shim_equals_direct_backward) pins it against the non-cached computation instead.  Call sites in the
reference: modules/training_utils.py:126-127 (cached forward), :141-144 (closure replay), :157/:247 (cat).
"""
from __future__ import annotations

import functools

import torch


class _RngSnapshot:
    """CPU (+ current CUDA device) RNG state captured at the no-grad forward and re-installed for the replay,
    so dropout masks of the second forward equal those of the first."""

    def __init__(self, *tensors):
        self.cpu = torch.get_rng_state()
        self.dev = None
        self.cuda_state = None
        for t in tensors:
            if isinstance(t, torch.Tensor) and t.is_cuda:
                self.dev = t.device
                self.cuda_state = torch.cuda.get_rng_state(t.device)
                break

    def __enter__(self):
        devices = [self.dev] if self.dev is not None else []
        self._fork = torch.random.fork_rng(devices=devices)
        self._fork.__enter__()
        torch.set_rng_state(self.cpu)
        if self.dev is not None:
            torch.cuda.set_rng_state(self.cuda_state, self.dev)
        return self

    def __exit__(self, *exc):
        return self._fork.__exit__(*exc)


def _tensors_in(obj):
    if isinstance(obj, torch.Tensor):
        yield obj
    elif isinstance(obj, (list, tuple)):
        for o in obj:
            yield from _tensors_in(o)
    elif hasattr(obj, "values"):
        for o in obj.values():
            yield from _tensors_in(o)


def cached(func):
    """Decorate a model call.  Returns (leaf, closure): `leaf` is the detached representation with
    requires_grad set (the loss is computed on it); `closure(leaf)` re-runs the call WITH grad under the
    captured RNG state and back-propagates <output, leaf.grad> into the model parameters."""

    @functools.wraps(func)
    def cache_func(*args, **kwargs):
        rng = _RngSnapshot(*_tensors_in(args), *_tensors_in(kwargs))
        with torch.no_grad():
            reps = func(*args, **kwargs)
        leaf = reps.detach().requires_grad_()

        def forward_backward_func(cache_reps):
            with rng:
                out = func(*args, **kwargs)
            surrogate = torch.dot(out.flatten(), cache_reps.grad.flatten().to(out.dtype))
            surrogate.backward()

        return leaf, forward_backward_func

    return cache_func


def _cat_if_list(x):
    if isinstance(x, list) and len(x) > 0 and isinstance(x[0], torch.Tensor):
        return torch.cat(x, dim=0)
    return x


def cat_input_tensor(func):
    """Concatenate every argument that is a non-empty list of tensors along dim 0 before the call."""

    @functools.wraps(func)
    def cat_f(*args, **kwargs):
        return func(*[_cat_if_list(a) for a in args], **{k: _cat_if_list(v) for k, v in kwargs.items()})

    return cat_f
