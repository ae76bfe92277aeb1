#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ops.py tests/test_gpu_unet.py -x -q -m gpu -k "upsample or unet" > gpurun_out/tests_up.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/tests_up.log
timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_b64.log 2>&1; grep '^{"metric"' gpurun_out/bench_b64.log | cut -c1-60,95-200
GSB_UP_BWD_WIN=0 timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_b64_old.log 2>&1; grep '^{"metric"' gpurun_out/bench_b64_old.log | cut -c1-60,95-200
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-graph"
timeout 600 $CMD > gpurun_out/plain_x.log 2>&1 &&
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches_x.csv $CMD > gpurun_out/ncu_list_x.log 2>&1
echo "ncu list rc=$?"
