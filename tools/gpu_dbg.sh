#!/bin/bash
mkdir -p gpurun_out
timeout 180 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 tools/peer_check.py > gpurun_out/peer_check.log 2>&1; echo "peer_check rc=$?"; grep "peer_check\|Error\|error" gpurun_out/peer_check.log | head -5
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29532 tools/dp_check.py > gpurun_out/dp_check_2.log 2>&1; echo "dp_check rc=$?"; grep "world\|worst" gpurun_out/dp_check_2.log
for CFG in "1:" "0:" "1:--no-sync-bn"; do
PEER=${CFG%%:*}; FLAGS=${CFG#*:}
GSB_PEER_ALLREDUCE=$PEER timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29544 bench.py --gpus 2 --steps 20 --warmup 3 --per-gpu-batch 8 $FLAGS > gpurun_out/bench_2x8.log 2>&1; echo "bench peer=$PEER $FLAGS rc=$?"
grep '^{"metric"' gpurun_out/bench_2x8.log | cut -c1-60,95-200
done
