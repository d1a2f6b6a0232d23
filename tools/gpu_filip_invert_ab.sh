#!/bin/bash
# warp-per-pair inversion against the CTA-per-pair kernel: gradient hashes must agree; then timing
mkdir -p gpurun_out
for inv in warp cta; do
  echo "== PCLIP_FILIP_INVERT=$inv" | tee -a gpurun_out/invert_ab.log
  PCLIP_FILIP_INVERT=$inv timeout 120 python tools/gpu_filip_gradhash.py 2>&1 | tail -6 | tee -a gpurun_out/invert_ab.log
  for mode in full ragged; do
    PCLIP_FILIP_INVERT=$inv timeout 120 python tools/gpu_filip_split.py $mode 2>&1 | grep -E "filip_bwd|total" | tee -a gpurun_out/invert_ab.log
  done
done
timeout 300 python -m pytest tests/test_gpu_filip.py -m gpu -q 2>&1 | tail -3
