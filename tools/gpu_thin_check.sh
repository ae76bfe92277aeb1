#!/bin/bash
mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_gpu_ops.py -q -k "thin" > gpurun_out/thin_tests.log 2>&1; echo "thin tests rc=$?"; tail -3 gpurun_out/thin_tests.log
timeout 200 python tools/bench_extra.py --only pix2pix --steps 10 2>/dev/null | tail -1 | cut -c1-170
GSB_THIN_PX=0 timeout 200 python tools/bench_extra.py --only pix2pix --steps 10 2>/dev/null | tail -1 | cut -c1-170
timeout 300 python -m pytest tests/test_gpu_pix2pix.py -q > gpurun_out/p2p_tests.log 2>&1; echo "pix2pix tests rc=$?"; tail -3 gpurun_out/p2p_tests.log
