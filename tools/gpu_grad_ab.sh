#!/bin/bash
# A/B of the CLIP gradient kernel instantiations (debug aid): parity tests then a short bench for each
mkdir -p gpurun_out
for epi in 8 16; do
  echo "== PCLIP_GRAD_EPI=$epi"
  PCLIP_GRAD_EPI=$epi timeout 400 python -m pytest tests/test_gpu_clip.py tests/test_gpu_sharded.py -m gpu -x -q 2>&1 | tail -4
  PCLIP_GRAD_EPI=$epi timeout 200 python bench.py --steps 10 --warmup 3 --no-extras 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['roofline'])"
done > gpurun_out/grad_ab.log 2>&1
cat gpurun_out/grad_ab.log
