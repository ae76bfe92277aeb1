#!/bin/bash
# r02v (8 GPUs): Pix2Pix at 1 / 4 / 8 ranks and GenSeg at 1 / 8 ranks with the steps replayed as CUDA graphs
mkdir -p gpurun_out
bash tools/gpu_scale.sh r02v_p2p "1 4 8" "pix2pix"
bash tools/gpu_scale.sh r02v_genseg "1 8" "genseg"
