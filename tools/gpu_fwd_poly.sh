#!/bin/bash
pr() { python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$1', d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['other_kernels_ms'])"; }
for K in 10 6 8 12 14 10 8; do PCLIP_FWD_POLY=$K timeout 200 python bench.py --steps 20 --warmup 5 --no-extras 2>&1 | tail -1 | pr "fwd_poly=$K"; done
