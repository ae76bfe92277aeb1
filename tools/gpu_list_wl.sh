#!/bin/bash
# launch list of ONE step of a secondary workload: tools/gpu_list_wl.sh <tag> <workload>  (eager launches; the plain run tells
# the launches per step L, ncu then skips 4 steps and records exactly L launches)
mkdir -p gpurun_out
TAG=${1:-wl}; WL=${2:-pix2pix}
CMD="python bench.py --workload $WL --steps 3 --warmup 3 --no-cpu-baseline --no-hbm-roofline --no-graph"
$CMD > gpurun_out/${TAG}_plain.json 2> gpurun_out/${TAG}_plain.err || { echo "plain run failed"; tail -5 gpurun_out/${TAG}_plain.err; exit 1; }
L=$(python -c "import json; d=json.loads(open('gpurun_out/${TAG}_plain.json').read().strip().splitlines()[-1]); print(d['gpu_launches']//d['steps'])")
echo "launches per step: $L"
ncu --metrics gpu__time_duration.sum --clock-control none -s $((L * 4)) -c $L --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu.log 2>&1
echo "list rc=$?"; python tools/launch_summary.py gpurun_out/${TAG}_launches.csv 45 | tee gpurun_out/${TAG}_launch_summary.txt
