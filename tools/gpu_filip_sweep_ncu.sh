#!/bin/bash
# ncu full capture of the single-sweep FILIP forward kernel (after a plain run of the same command)
mkdir -p gpurun_out
run_filip="python tools/gpu_filip_prof.py"
$run_filip > gpurun_out/p_plain_filip.log 2>&1 &&
ncu --clock-control none --set full --import-source on -k regex:tc_filip_both_kernel -s 2 -c 1 -f -o gpurun_out/r02_filip_both $run_filip > gpurun_out/ncu_filip_both.log 2>&1
tail -3 gpurun_out/p_plain_filip.log gpurun_out/ncu_filip_both.log
ls -la gpurun_out/*.ncu-rep
