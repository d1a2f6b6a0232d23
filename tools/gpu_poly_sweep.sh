#!/bin/bash
# exp split sweeps (debug aid): share of exps on the FMA pipes in the gradient (build variants) and forward (env) kernels
LIB=protein-clip_b200/pclip/libpclip.so
cp $LIB /tmp/libpclip_orig.so
pr() { python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$1', d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['other_kernels_ms'])"; }
for K in 0 4 12 16; do cp build/libpclip_bp$K.so $LIB; timeout 200 python bench.py --steps 10 --warmup 3 --no-extras 2>&1 | tail -1 | pr "bwd_poly=$K"; done
cp /tmp/libpclip_orig.so $LIB
for K in 6 10 14 18; do PCLIP_FWD_POLY=$K timeout 200 python bench.py --steps 10 --warmup 3 --no-extras 2>&1 | tail -1 | pr "bwd_poly=8 fwd_poly=$K"; done
