#!/bin/bash
mkdir -p gpurun_out
GSB_THIN_SPX=1 timeout 200 python -m pytest tests/test_gpu_ops.py -q -k "thin" > gpurun_out/spx_tests.log 2>&1; echo "thin tests (SPX=1) rc=$?"; tail -3 gpurun_out/spx_tests.log
GSB_THIN_SPX=1 timeout 300 python -m pytest tests/test_gpu_pix2pix.py -q > gpurun_out/spx_p2p_tests.log 2>&1; echo "pix2pix tests (SPX=1) rc=$?"; tail -3 gpurun_out/spx_p2p_tests.log
GSB_THIN_SPX=1 timeout 200 python tools/bench_extra.py --only pix2pix --steps 10 2>/dev/null | tail -1 | cut -c1-170
GSB_THIN_SPX=0 timeout 200 python tools/bench_extra.py --only pix2pix --steps 10 2>/dev/null | tail -1 | cut -c1-170
bash tools/gpu_list_final.sh
