#!/bin/bash
# launch list of the timed bench steps (ncu, per-launch durations) for N images per GPU: tools/gpu_list.sh <tag> [per-gpu-batch]
mkdir -p gpurun_out
TAG=${1:-list}; B=${2:-0}
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --per-gpu-batch $B"
$CMD > gpurun_out/${TAG}_plain.json 2> gpurun_out/${TAG}_plain.err || { echo "plain run failed"; tail -5 gpurun_out/${TAG}_plain.err; exit 1; }
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu.log 2>&1
echo "list rc=$?"; python tools/launch_summary.py gpurun_out/${TAG}_launches.csv 40 | tee gpurun_out/${TAG}_launch_summary.txt
