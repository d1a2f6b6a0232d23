"""Minimal `grad_cache` package: the reference imports `grad_cache.functional` (modules/training_utils.py:2) from
luyug/GradCache, an un-vendored submodule (pinned 33695437) that is absent from the reference checkout."""
from . import functional  # noqa: F401
