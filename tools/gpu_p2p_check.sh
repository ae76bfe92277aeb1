#!/bin/bash
mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_gpu_ops.py -q -k "stem or thin or edge" > gpurun_out/stem_tests.log 2>&1; echo "stem tests rc=$?"; tail -3 gpurun_out/stem_tests.log
timeout 300 python -m pytest tests/test_gpu_pix2pix.py -q > gpurun_out/p2p_tests.log 2>&1; echo "pix2pix tests rc=$?"; tail -3 gpurun_out/p2p_tests.log
timeout 200 python tools/bench_extra.py --only pix2pix --steps 10 > gpurun_out/p2p.jsonl 2> gpurun_out/p2p.err; cat gpurun_out/p2p.jsonl
GSB_STEM_WGRAD_FAST42=0 timeout 200 python tools/bench_extra.py --only pix2pix --steps 10 2>/dev/null | tail -1
