#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/b1_check.py > gpurun_out/b1_check.log 2>&1; echo "b1 rc=$?"; grep "n=" gpurun_out/b1_check.log
for GB in 2 8; do
DP_GB=$GB timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2951$GB tools/dp_check.py > gpurun_out/dp_check_gb$GB.log 2>&1; echo "dp_check rc=$?"; grep "world\|worst" gpurun_out/dp_check_gb$GB.log
done
timeout 900 python -m pytest tests/test_gpu_ops.py -x -q -m gpu -k "multi or batched or thin or optim" > gpurun_out/newtests.log 2>&1; echo "tests rc=$?"; tail -5 gpurun_out/newtests.log
