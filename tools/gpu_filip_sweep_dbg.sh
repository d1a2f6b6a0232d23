#!/bin/bash
# timing experiments on the single-sweep FILIP forward: LDTM probe, then the kernel without epilogue arithmetic / loads
mkdir -p gpurun_out
timeout 60 tools/bin/probe_ldtm 2>&1 | tee gpurun_out/probe_ldtm.log
for dbg in 0 1 2; do
  echo "== PCLIP_FILIP_DBG=$dbg full" | tee -a gpurun_out/sweep_dbg.log
  PCLIP_FILIP_DBG=$dbg timeout 120 python tools/gpu_filip_split.py full 2>&1 | grep filip_fwd | tee -a gpurun_out/sweep_dbg.log
done
