#!/bin/bash
# quick GPU check: conv op tests + U-Net tests + short bench.  Usage: tools/gpu_quick.sh [tag] [pytest -k expr]
TAG=${1:-q}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ops.py -x -q -k "${2:-conv}" > gpurun_out/ops_$TAG.log 2>&1; echo "ops rc=$?"; tail -15 gpurun_out/ops_$TAG.log
timeout 900 python -m pytest tests/test_gpu_unet.py -x -q > gpurun_out/unet_$TAG.log 2>&1; echo "unet rc=$?"; tail -5 gpurun_out/unet_$TAG.log
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_$TAG.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/bench_$TAG.log
