#!/bin/bash
# launch list of one bench step (ncu, per-launch durations) at N images per GPU: tools/gpu_list.sh <tag> [per-gpu-batch]
# eager launches (--no-graph: same kernels as the replayed graph); 6 steps are skipped (3 warm-up + 3 settling), then ~2.3 steps
# are recorded and launch_summary.py --one-step cuts out exactly one.
mkdir -p gpurun_out
TAG=${1:-list}; B=${2:-0}
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-hbm-roofline --no-graph --per-gpu-batch $B"
$CMD > gpurun_out/${TAG}_plain.json 2> gpurun_out/${TAG}_plain.err || { echo "plain run failed"; tail -5 gpurun_out/${TAG}_plain.err; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -s 1040 -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu.log 2>&1
echo "list rc=$?"; python tools/launch_summary.py gpurun_out/${TAG}_launches.csv 40 --one-step | tee gpurun_out/${TAG}_launch_summary.txt
