#!/bin/bash
# single-sweep FILIP forward: share of the keys made on the FMA pipe (PCLIP_FILIP_FMAKEYS of 32), parity first
mkdir -p gpurun_out
timeout 200 python tools/gpu_filip_sweep_check.py 2>&1 | tail -3
for fma in 0 8 16 24 32; do
  echo "== fma keys $fma" | tee -a gpurun_out/sweep_fma.log
  PCLIP_FILIP_FMAKEYS=$fma timeout 120 python tools/gpu_filip_split.py full 2>&1 | grep filip_fwd | tee -a gpurun_out/sweep_fma.log
  PCLIP_FILIP_FMAKEYS=$fma timeout 120 python tools/gpu_filip_split.py ragged 2>&1 | grep filip_fwd | tee -a gpurun_out/sweep_fma.log
done
PCLIP_FILIP_FMAKEYS=16 timeout 200 python tools/gpu_filip_sweep_check.py 2>&1 | tail -2
