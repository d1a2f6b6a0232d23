#!/bin/bash
# mbarrier wait loop: suspend-time hint A/B on the CLIP step and the FILIP forward (libraries built with -DPCLIP_MBAR_HINT_NS=...)
mkdir -p gpurun_out
cp protein-clip_b200/pclip/libpclip.so /tmp/libpclip_orig.so
for h in 200 4000 20000 200 4000 20000; do
  cp tools/bin/ab/libpclip_h$h.so protein-clip_b200/pclip/libpclip.so
  echo "== hint $h ns" | tee -a gpurun_out/wait_ab.log
  timeout 200 python bench.py --steps 10 --warmup 3 --no-extras 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('clip ms', d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['other_kernels_ms'])" | tee -a gpurun_out/wait_ab.log
  timeout 120 python tools/gpu_filip_split.py full 2>&1 | grep "filip_fwd" | tee -a gpurun_out/wait_ab.log
done
cp /tmp/libpclip_orig.so protein-clip_b200/pclip/libpclip.so
