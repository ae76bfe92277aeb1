#!/bin/bash
# r02c: centered fp32 BatchNorm, rank-invariant dropout, peer watchdog; launch list of the bench step at 64 and 8 images per GPU
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --deselect tests/test_gpu_protocol.py -k "not production" > gpurun_out/r02c_suite.log 2>&1; echo "suite rc=$?"; tail -15 gpurun_out/r02c_suite.log
timeout 900 python -m pytest tests/test_gpu_protocol.py -m gpu -q -s > gpurun_out/r02c_protocol.log 2>&1; echo "protocol rc=$?"; tail -8 gpurun_out/r02c_protocol.log
grep -n "trajectory\|pix2pix fp32\|median" gpurun_out/r02c_protocol.log | cut -c1-330
bash tools/gpu_list.sh r02c_b64 0
bash tools/gpu_list.sh r02c_b8 8
tail -c 400 gpurun_out/r02c_b64_plain.json
