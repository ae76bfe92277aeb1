#!/bin/bash
# r02i: no-alloc entry point test (alone in a fresh process, then after the whole suite), CPU-visible checks
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x > gpurun_out/r02i_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"; tail -5 gpurun_out/r02i_noalloc_alone.log
timeout 2400 python -m pytest tests -m gpu -q -x > gpurun_out/r02i_suite.log 2>&1; echo "suite rc=$?"; tail -6 gpurun_out/r02i_suite.log
