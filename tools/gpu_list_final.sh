#!/bin/bash
# ncu launch list of one bench step (final state of the round); run only after the same command exited 0 without ncu.
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-graph"
timeout 200 $CMD > gpurun_out/plain_final.log 2>&1 &&
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches_final.csv $CMD > gpurun_out/ncu_list_final.log 2>&1
echo rc=$?
