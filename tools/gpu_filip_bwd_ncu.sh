#!/bin/bash
mkdir -p gpurun_out
NCU="ncu --clock-control none"
run_filip="python tools/gpu_filip_prof.py"
$run_filip > gpurun_out/p_plain_filip.log 2>&1 &&
$NCU --set full --import-source on -k regex:filip_invert_warp_kernel -s 2 -c 1 -f -o gpurun_out/r02b_filip_invert_warp $run_filip > /dev/null 2>&1
$NCU --set full --import-source on -k regex:filip_gather_grad_kernel -s 2 -c 1 -f -o gpurun_out/r02b_filip_gather $run_filip > /dev/null 2>&1
tail -1 gpurun_out/p_plain_filip.log; ls -la gpurun_out/*.ncu-rep
