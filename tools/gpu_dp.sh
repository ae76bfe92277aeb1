#!/bin/bash
# multi-GPU check: N-rank == 1-rank parity (sync-BN + global Dice), then the bench at N ranks and small-batch host overhead
N=${1:-2}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_pix2pix.py -x -q > gpurun_out/p2p_dp.log 2>&1; echo "pix2pix rc=$?"; tail -2 gpurun_out/p2p_dp.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tools/dp_check.py > gpurun_out/dp_check_$N.log 2>&1; echo "dp_check rc=$?"; grep world gpurun_out/dp_check_$N.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_dp$N.log 2>&1; echo "bench$N rc=$?"; tail -1 gpurun_out/bench_dp$N.log | cut -c1-400
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_dp1.log 2>&1; tail -1 gpurun_out/bench_dp1.log | cut -c1-330
timeout 600 python bench.py --per-gpu-batch 8 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_b8.log 2>&1; tail -1 gpurun_out/bench_b8.log | cut -c1-330
