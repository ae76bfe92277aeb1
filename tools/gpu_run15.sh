#!/bin/bash
# r02o (8 GPUs): scaling lines of the headline at 1/2/4/8 ranks and of Pix2Pix / GenSeg at 1 and 8 ranks, N-rank == 1-rank check at 8
mkdir -p gpurun_out
nvidia-smi -L | wc -l
bash tools/gpu_scale.sh r02o_unet "1 2 4 8" "unet"
bash tools/gpu_scale.sh r02o_gan "1 8" "pix2pix genseg"
DP_OVERLAP=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29590 tools/dp_check.py 2>&1 | grep -v "initialize\|Warning\|warn" | tee gpurun_out/r02o_dp_check_8gpu.log | tail -8
