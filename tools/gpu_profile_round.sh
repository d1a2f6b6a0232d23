#!/bin/bash
# One GPU-box visit: ncu launch lists and full captures of the dominant kernels of every workload (round 2).
# Every ncu run follows a plain run of the same command that exited 0.
mkdir -p gpurun_out
NCU="ncu --clock-control none"
run_clip="python bench.py --steps 2 --warmup 1 --no-extras"
$run_clip > gpurun_out/p_plain_clip.log 2>&1 &&
$NCU --metrics gpu__time_duration.sum -c 300 --csv --log-file gpurun_out/r02_clip_b65536_launches.csv $run_clip > /dev/null 2>&1
$NCU --set full --import-source on -k regex:tc3_grad_kernel -s 1 -c 1 -f -o gpurun_out/r02_tc3_grad $run_clip > /dev/null 2>&1
$NCU --set full --import-source on -k regex:tc2_sumexp_kernel -s 1 -c 1 -f -o gpurun_out/r02_tc2_sumexp $run_clip > /dev/null 2>&1
run_filip="python tools/gpu_filip_prof.py"
$run_filip > gpurun_out/p_plain_filip.log 2>&1 &&
$NCU --metrics gpu__time_duration.sum -c 200 --csv --log-file gpurun_out/r02_filip_b256_launches.csv $run_filip > /dev/null 2>&1
$NCU --set full --import-source on -k regex:tc_filip_max_kernel -s 2 -c 1 -f -o gpurun_out/r02_filip_max $run_filip > /dev/null 2>&1
$NCU --set full --import-source on -k regex:filip_gather_grad_kernel -s 2 -c 1 -f -o gpurun_out/r02_filip_gather $run_filip > /dev/null 2>&1
$NCU --set full --import-source on -k regex:filip_invert_kernel -s 2 -c 1 -f -o gpurun_out/r02_filip_invert $run_filip > /dev/null 2>&1
python tools/gpu_filip_prof.py ragged > gpurun_out/p_plain_filip_r.log 2>&1 &&
$NCU --metrics gpu__time_duration.sum -c 200 --csv --log-file gpurun_out/r02_filip_b256_ragged_launches.csv python tools/gpu_filip_prof.py ragged > /dev/null 2>&1
run_topk="python tools/gpu_topk_prof.py 37888"
$run_topk > gpurun_out/p_plain_topk.log 2>&1 &&
$NCU --set full --import-source on -k regex:tc_topk_kernel -s 1 -c 1 -f -o gpurun_out/r02_tc_topk $run_topk > /dev/null 2>&1
run_small="python tools/gpu_topk_small.py 128"
$run_small > gpurun_out/p_plain_small.log 2>&1 &&
$NCU --set full -k regex:tc_topk_kernel -s 3 -c 1 -f -o gpurun_out/r02_tc_topk_q128 $run_small > /dev/null 2>&1
run_gc="python tools/gpu_gc_prof.py"
$run_gc > gpurun_out/p_plain_gc.log 2>&1 &&
$NCU --set full -k regex:clip_small_fused_kernel -s 5 -c 1 -f -o gpurun_out/r02_clip_small $run_gc > /dev/null 2>&1
cat gpurun_out/p_plain_*.log | tail -20
ls -la gpurun_out/*.ncu-rep
