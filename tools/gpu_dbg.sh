#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_unet.py -x -q -m gpu > gpurun_out/tests_unet.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/tests_unet.log
for OV in 0 1; do
DP_OVERLAP=$OV timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2953$OV tools/dp_check.py > gpurun_out/dp_check_ov$OV.log 2>&1; echo "dp_check overlap=$OV rc=$?"; grep "world\|Error" gpurun_out/dp_check_ov$OV.log | head
done
for FLAGS in "" "--no-overlap" "--no-graph" "--no-sync-bn"; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29544 bench.py --gpus 2 --steps 20 --warmup 3 --per-gpu-batch 8 $FLAGS > gpurun_out/bench_2x8.log 2>&1; echo "bench [$FLAGS] rc=$?"
grep '^{"metric"' gpurun_out/bench_2x8.log | cut -c1-60,95-200; grep -i "error\|Traceback" gpurun_out/bench_2x8.log | head -3
done
for FLAGS in "" "--no-overlap"; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29545 bench.py --gpus 2 --steps 20 --warmup 3 $FLAGS > gpurun_out/bench_2x32.log 2>&1; echo "bench 2x32 [$FLAGS] rc=$?"
grep '^{"metric"' gpurun_out/bench_2x32.log | cut -c1-60,95-200
done
