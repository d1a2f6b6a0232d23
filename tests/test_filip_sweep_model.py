"""The index arithmetic of the single-sweep FILIP forward (protein-clip_b200/csrc/filip_tc1.cu) on the CPU: a lane-level
numpy model of the tcgen05.ld.16x256b register layout, the shared 5-bit key code, the per-tile row update, the
shared-memory column exchange and the finaliser must reproduce argmax / max of a score matrix in both directions
(ref: modules/models.py:332-351, amax over the column tokens and over the row tokens).  The kernel itself is held to the
oracle by tests/test_gpu_filip.py; this pins the layout the kernel's comments describe."""
import os
import sys

import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tools"))


@pytest.mark.parametrize("lx,ly", [(77, 131), (129, 33), (1, 1), (40, 260)])
def test_lane_model_reproduces_row_and_column_argmax(lx, ly):
    import filip_sweep_model as m
    m.run(lx, ly)   # asserts agreement of both argmax directions and the 2^-19 value bound
