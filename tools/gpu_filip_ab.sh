#!/bin/bash
# A/B of the FILIP forward y-split (PCLIP_FILIP_YSPLIT) on full-length and ragged batches, after the parity tests
timeout 300 python -m pytest tests/test_gpu_filip.py -m gpu -x -q 2>&1 | tail -3
for ys in 0 1 2 4 8; do
  for mode in full ragged; do
    echo "== ysplit=$ys $mode"
    PCLIP_FILIP_YSPLIT=$ys timeout 120 python tools/gpu_filip_split.py $mode 2>&1 | tail -4
  done
done
