#!/bin/bash
# forward-kernel parity print + tuning sweep (debug aid)
mkdir -p gpurun_out
timeout 300 python tools/gpu_debug_sumexp.py > gpurun_out/dbg_sumexp.log 2>&1; echo "exit $?" >> gpurun_out/dbg_sumexp.log
for p in 0 6 10 14 18; do echo "== PCLIP_FWD_POLY=$p"; QUICK=1 PCLIP_FWD_POLY=$p timeout 200 python tools/gpu_debug_sumexp.py 2>&1 | grep "single pass"; done > gpurun_out/fwd_sweep.log 2>&1
tail -18 gpurun_out/dbg_sumexp.log; cat gpurun_out/fwd_sweep.log
