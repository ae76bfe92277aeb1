#!/bin/bash
# r02j (2 GPUs): weight gradients on the side stream by default; whole suite incl. the 2-rank test, dp_check (overlap on/off),
# 1- and 2-GPU bench lines of every workload
mkdir -p gpurun_out
nvidia-smi -L
timeout 1800 python -m pytest tests -m gpu -q -x > gpurun_out/r02j_suite.log 2>&1; echo "suite rc=$?"; tail -5 gpurun_out/r02j_suite.log
for OV in 0 1; do
DP_OVERLAP=$OV timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2957$OV tools/dp_check.py 2>&1 | grep -v "initialize\|Warning\|warn" | tee gpurun_out/r02j_dp_check_overlap$OV.log | tail -10
done
bash tools/gpu_scale.sh r02j "1 2" "unet pix2pix genseg"
