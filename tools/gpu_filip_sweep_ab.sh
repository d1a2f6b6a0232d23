#!/bin/bash
# single-sweep FILIP forward: parity against the fp32 path, the FILIP test file, then A/B timing against two passes
mkdir -p gpurun_out
timeout 200 python tools/gpu_filip_sweep_check.py > gpurun_out/sweep_check.log 2>&1; echo "check rc=$?" >> gpurun_out/sweep_check.log
tail -30 gpurun_out/sweep_check.log
timeout 500 python -m pytest tests/test_gpu_filip.py -m gpu -q 2>&1 | tail -25 > gpurun_out/sweep_tests.log
tail -25 gpurun_out/sweep_tests.log
for mode in full ragged; do
  for fwd in 0 2; do
    echo "== PCLIP_FILIP_FWD=$fwd $mode" | tee -a gpurun_out/sweep_ab.log
    PCLIP_FILIP_FWD=$fwd timeout 120 python tools/gpu_filip_split.py $mode 2>&1 | tail -7 | tee -a gpurun_out/sweep_ab.log
  done
done
for ys in 2 8 16; do
  echo "== single sweep ysplit=$ys full" | tee -a gpurun_out/sweep_ab.log
  PCLIP_FILIP_YSPLIT=$ys timeout 120 python tools/gpu_filip_split.py full 2>&1 | grep filip_fwd | tee -a gpurun_out/sweep_ab.log
done
