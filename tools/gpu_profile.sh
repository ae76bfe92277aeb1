#!/bin/bash
# One GPU-box call: GPU parity tests, smoke, bench, then (only if the plain run exited 0) the ncu launch list of
# the same bench command and one full capture of the conv kernels.  Usage: tools/gpu_profile.sh [tag]
TAG=${1:-r01}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
if [ -z "$SKIP_TESTS" ]; then
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?"
tail -6 gpurun_out/pytest_gpu_$TAG.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_$TAG.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke_$TAG.log
fi
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_$TAG.log 2>&1; echo "bench rc=$?"; tail -2 gpurun_out/bench_$TAG.log
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_$TAG.log 2>&1; echo "ref rc=$?"; tail -1 gpurun_out/bench_ref_$TAG.log
if [ -z "$SKIP_NCU" ]; then
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-graph"
timeout 600 $CMD > gpurun_out/plain_$TAG.log 2>&1 &&
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv \
    --log-file gpurun_out/launches_$TAG.csv $CMD > gpurun_out/ncu_list_$TAG.log 2>&1
echo "ncu list rc=$?"
for K in ${NCU_KERNELS:-conv_tc_halo_kernel}; do
timeout 600 $CMD > gpurun_out/plain2_$TAG.log 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on --profile-from-start off \
    -k regex:"$K" -s ${NCU_SKIP:-0} -c ${NCU_COUNT:-4} -f -o gpurun_out/prof_${TAG}_$K $CMD > gpurun_out/ncu_full_${TAG}_$K.log 2>&1
echo "ncu full $K rc=$?"
ncu -i gpurun_out/prof_${TAG}_$K.ncu-rep --page raw --csv > gpurun_out/prof_${TAG}_$K.raw.csv 2>/dev/null
done
fi
du -sm gpurun_out | tail -1
# gpurun copies back at most 64 MiB: drop the largest reports first (the raw CSV summaries stay)
while [ $(du -sm gpurun_out | cut -f1) -gt 60 ]; do
  big=$(ls -S gpurun_out/*.ncu-rep 2>/dev/null | head -1); [ -z "$big" ] && break; echo "dropping $big"; rm -f "$big"
done
