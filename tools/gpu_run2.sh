#!/bin/bash
# r02b: deterministic reductions -- whole GPU suite, then bench
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --deselect tests/test_gpu_protocol.py > gpurun_out/r02b_suite.log 2>&1; echo "suite rc=$?"; tail -30 gpurun_out/r02b_suite.log
timeout 900 python -m pytest tests/test_gpu_protocol.py -m gpu -q -s > gpurun_out/r02b_protocol.log 2>&1; echo "protocol rc=$?"; tail -12 gpurun_out/r02b_protocol.log
timeout 300 python bench.py > gpurun_out/r02b_bench.json 2> gpurun_out/r02b_bench.err; echo "bench rc=$?"; tail -c 900 gpurun_out/r02b_bench.json; tail -3 gpurun_out/r02b_bench.err
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
