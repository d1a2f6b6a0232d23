#!/bin/bash
# A/B of the two CLIP gradient formulations (debug aid): parity tests then a short bench for each
mkdir -p gpurun_out
for form in t r; do
  echo "== PCLIP_GRAD_FORM=$form"
  PCLIP_GRAD_FORM=$form timeout 400 python -m pytest tests/test_gpu_clip.py tests/test_gpu_sharded.py -m gpu -x -q 2>&1 | tail -4
  PCLIP_GRAD_FORM=$form timeout 200 python bench.py --steps 10 --warmup 3 --no-extras 2>&1 | tail -1
done > gpurun_out/grad_ab.log 2>&1
PCLIP_BWD_TRACE=gpurun_out/bwd_trace_t.txt timeout 120 python tools/gpu_bwd_trace.py > /dev/null 2>&1
cat gpurun_out/grad_ab.log; head -30 gpurun_out/bwd_trace_t.txt
