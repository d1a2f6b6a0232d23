#!/bin/bash
for rpw in 4 8; do
  echo "== PCLIP_FILIP_RPW=$rpw"
  PCLIP_FILIP_RPW=$rpw timeout 120 python tools/gpu_filip_gradhash.py 2>&1 | grep grads | head -3
  for mode in full ragged; do PCLIP_FILIP_RPW=$rpw timeout 120 python tools/gpu_filip_split.py $mode 2>&1 | grep -E "filip_bwd"; done
done
