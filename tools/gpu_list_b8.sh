#!/bin/bash
mkdir -p gpurun_out
CMD="python bench.py --per-gpu-batch 8 --steps 1 --warmup 3 --no-cpu-baseline --no-graph"
timeout 600 $CMD > gpurun_out/plain_b8.log 2>&1 &&
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches_b8.csv $CMD > gpurun_out/ncu_list_b8.log 2>&1
echo rc=$?
