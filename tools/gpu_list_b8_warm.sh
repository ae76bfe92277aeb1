#!/bin/bash
# Launch list at 8 images per GPU with warm caches (--cache-control none): closer to the in-graph durations than the
# default flush-between-kernels list.  Also the grid sizes, to see which launches under-fill the GPU.
mkdir -p gpurun_out
CMD="python bench.py --per-gpu-batch 8 --steps 1 --warmup 3 --no-cpu-baseline --no-graph"
timeout 300 $CMD > gpurun_out/plain_b8.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum,launch__grid_size,launch__block_size --clock-control none --cache-control none --profile-from-start off --csv --log-file gpurun_out/launches_b8_warm.csv $CMD > gpurun_out/ncu_list_b8.log 2>&1
echo rc=$?
