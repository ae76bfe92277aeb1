#!/bin/bash
# 2-GPU regression after the optimiser / launch-helper changes: 2-rank GPU tests, N-rank == 1-rank check, bench at 2 ranks.
mkdir -p gpurun_out
timeout 240 python -m pytest tests/test_gpu_parallel.py -q > gpurun_out/g2_tests.log 2>&1; echo "parallel tests rc=$?"; tail -3 gpurun_out/g2_tests.log
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/dp_check.py > gpurun_out/g2_dp_check.log 2>&1; echo "dp_check rc=$?"; grep -i "world\|loss\|grad" gpurun_out/g2_dp_check.log | head -8
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/g2_bench.json 2> gpurun_out/g2_bench.err; echo "bench2 rc=$?"; tail -1 gpurun_out/g2_bench.json | cut -c1-500
