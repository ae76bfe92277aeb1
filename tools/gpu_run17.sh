#!/bin/bash
# r02q: vectorised weight-gradient epilogue for transposed convolutions, thin_tc on by default: whole suite + workloads + lists
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q -x > gpurun_out/r02q_suite.log 2>&1; echo "suite rc=$?"; tail -4 gpurun_out/r02q_suite.log | cut -c1-300
bash tools/gpu_scale.sh r02q "1" "unet pix2pix genseg"
bash tools/gpu_list_wl.sh r02q_p2p pix2pix | head -16
