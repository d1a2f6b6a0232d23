#!/bin/bash
for ys in 0 4 8 16 32; do
  echo "== ysplit=$ys"
  PCLIP_FILIP_YSPLIT=$ys timeout 120 python tools/gpu_filip_split.py ragged 2>&1 | grep filip_fwd
  PCLIP_FILIP_YSPLIT=$ys timeout 120 python tools/gpu_filip_split.py full 2>&1 | grep filip_fwd
done
