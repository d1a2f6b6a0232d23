#!/bin/bash
# A/B of the two CLIP gradient formulations (debug aid): parity tests then a short bench for each
mkdir -p gpurun_out
for form in t r; do
  echo "== PCLIP_GRAD_FORM=$form"
  PCLIP_GRAD_FORM=$form timeout 400 python -m pytest tests/test_gpu_clip.py tests/test_gpu_sharded.py -m gpu -x -q 2>&1 | tail -2
  for i in 1 2; do PCLIP_GRAD_FORM=$form timeout 200 python bench.py --steps 10 --warmup 3 --no-extras 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['other_kernels_ms'])"; done
done > gpurun_out/grad_ab.log 2>&1
cat gpurun_out/grad_ab.log
