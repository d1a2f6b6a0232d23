#!/bin/bash
# CLIP profile visit: launch list of the headline command, full captures of the gradient and forward kernels
mkdir -p gpurun_out
NCU="ncu --clock-control none"
run_clip="python bench.py --steps 2 --warmup 1 --no-extras"
$run_clip > gpurun_out/p_plain_clip.log 2>&1 &&
$NCU --metrics gpu__time_duration.sum -c 300 --csv --log-file gpurun_out/r02b_clip_b65536_launches.csv $run_clip > /dev/null 2>&1
$NCU --set full --import-source on -k regex:tc3_grad_kernel -s 1 -c 1 -f -o gpurun_out/r02b_tc3_grad $run_clip > /dev/null 2>&1
$NCU --set full --import-source on -k regex:tc2_sumexp_kernel -s 1 -c 1 -f -o gpurun_out/r02b_tc2_sumexp $run_clip > /dev/null 2>&1
tail -c 600 gpurun_out/p_plain_clip.log; ls -la gpurun_out/r02b_tc*.ncu-rep
