#!/bin/bash
# ncu --set full of selected kernels of one eager bench step: tools/gpu_ncu_full.sh <tag> <kernel regex> [count] [per-gpu-batch]
# (run only after the same command has exited 0 without ncu; the .ncu-rep comes back in gpurun_out/ and is read on the
# build machine with `ncu -i ... --page raw --csv`)
mkdir -p gpurun_out
TAG=${1:-full}; RE=${2:-stem_tc_fwd}; CNT=${3:-4}; B=${4:-0}
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-hbm-roofline --no-graph --per-gpu-batch $B"
$CMD > gpurun_out/${TAG}_plain.json 2> gpurun_out/${TAG}_plain.err || { echo "plain run failed"; tail -5 gpurun_out/${TAG}_plain.err; exit 1; }
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:$RE" -s $((CNT * 3)) -c $CNT -f -o gpurun_out/${TAG} $CMD > gpurun_out/${TAG}_ncu.log 2>&1
echo "ncu rc=$?"; ls -la gpurun_out/${TAG}.ncu-rep
ncu -i gpurun_out/${TAG}.ncu-rep --page raw --csv > gpurun_out/${TAG}.raw.csv 2>/dev/null; wc -c gpurun_out/${TAG}.raw.csv
