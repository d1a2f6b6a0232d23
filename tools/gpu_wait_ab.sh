#!/bin/bash
# does the mbarrier wait loop (suspend hint) cost the CLIP / two-pass FILIP kernels anything?  (library built twice)
mkdir -p gpurun_out
for lib in hint nohint; do
  cp build/ab/libpclip_$lib.so protein-clip_b200/pclip/libpclip.so
  echo "== $lib" | tee -a gpurun_out/wait_ab.log
  timeout 200 python bench.py --steps 10 --warmup 3 --no-extras 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('clip ms', d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['other_kernels_ms'])" | tee -a gpurun_out/wait_ab.log
  for fwd in 0 2; do PCLIP_FILIP_FWD=$fwd timeout 120 python tools/gpu_filip_split.py full 2>&1 | grep "filip_fwd" | tee -a gpurun_out/wait_ab.log; done
done
