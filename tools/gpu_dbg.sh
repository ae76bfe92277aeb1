#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/tests_all.log 2>&1; echo "tests rc=$?"; tail -4 gpurun_out/tests_all.log
timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_b64.log 2>&1; grep '^{"metric"' gpurun_out/bench_b64.log | cut -c1-60,95-200
GSB_BN_BWD_FAST2=0 timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_b64_old.log 2>&1; grep '^{"metric"' gpurun_out/bench_b64_old.log | cut -c1-60,95-200
timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --per-gpu-batch 8 > gpurun_out/bench_b8.log 2>&1; grep '^{"metric"' gpurun_out/bench_b8.log | cut -c1-60,95-200
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-graph"
timeout 600 $CMD > gpurun_out/plain_x.log 2>&1 &&
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches_x.csv $CMD > gpurun_out/ncu_list_x.log 2>&1
echo "ncu list rc=$?"
