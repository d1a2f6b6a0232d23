#!/bin/bash
# final FILIP profile visit of round 2: launch lists (full, ragged), full capture of the warp-per-pair inversion
mkdir -p gpurun_out
NCU="ncu --clock-control none"
run_filip="python tools/gpu_filip_prof.py"
$run_filip > gpurun_out/p_plain_filip.log 2>&1 &&
$NCU --metrics gpu__time_duration.sum -c 200 --csv --log-file gpurun_out/r02b_filip_b256_launches.csv $run_filip > /dev/null 2>&1
$NCU --set full --import-source on -k regex:filip_invert_warp_kernel -s 2 -c 1 -f -o gpurun_out/r02b_filip_invert_warp $run_filip > /dev/null 2>&1
python tools/gpu_filip_prof.py ragged > gpurun_out/p_plain_filip_r.log 2>&1 &&
$NCU --metrics gpu__time_duration.sum -c 200 --csv --log-file gpurun_out/r02b_filip_b256_ragged_launches.csv python tools/gpu_filip_prof.py ragged > /dev/null 2>&1
cat gpurun_out/p_plain_filip.log gpurun_out/p_plain_filip_r.log | tail -4
