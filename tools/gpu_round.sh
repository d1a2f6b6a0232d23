#!/bin/bash
# One GPU-box visit: parity tests, smoke, bench (both arms), ncu launch list and one full capture of the top kernel.
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/tests.log 2>&1; echo "tests exit $?" >> gpurun_out/tests.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke.log
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?" >> gpurun_out/bench.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2>&1
KPAT=${KPAT:-tc3_grad_kernel}
python bench.py --steps 2 --warmup 1 --no-extras > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 1 --no-extras > gpurun_out/ncu1.log 2>&1
python bench.py --steps 2 --warmup 1 --no-extras > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$KPAT -s 1 -c 1 -f -o gpurun_out/prof_top \
    python bench.py --steps 2 --warmup 1 --no-extras > gpurun_out/ncu2.log 2>&1
KPAT2=${KPAT2:-tc2_sumexp_kernel}
ncu --set full --clock-control none --import-source on -k regex:$KPAT2 -s 1 -c 1 -f -o gpurun_out/prof_second \
    python bench.py --steps 2 --warmup 1 --no-extras > gpurun_out/ncu3.log 2>&1
tail -3 gpurun_out/tests.log; tail -2 gpurun_out/smoke.log; cat gpurun_out/bench.json; tail -2 gpurun_out/bench.err
