import os
import sys

import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(REPO, "protein-clip_b200")
for p in (REPO, HOST):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(REPO, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    import torch

    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def golden():
    import numpy as np
    import torch

    def load(name, dtype=None):
        z = np.load(os.path.join(GOLDEN, name + ".npz"))
        out = {}
        for k in z.files:
            t = torch.from_numpy(z[k])
            if dtype is not None and t.is_floating_point():
                t = t.to(dtype)
            out[k] = t
        return out

    return load
