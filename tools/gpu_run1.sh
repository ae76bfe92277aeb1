#!/bin/bash
# r02a: new parity tests first (verbose, not -x), then the whole suite, bench, smoke
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_protocol.py -m gpu -q -s > gpurun_out/r02a_protocol.log 2>&1; echo "protocol rc=$?"; tail -15 gpurun_out/r02a_protocol.log
timeout 900 python -m pytest tests/test_gpu_ops.py -m gpu -q -k production -s > gpurun_out/r02a_prod.log 2>&1; echo "prod rc=$?"; tail -5 gpurun_out/r02a_prod.log
timeout 600 python -m pytest tests -m gpu -q --deselect tests/test_gpu_protocol.py -k "not production" > gpurun_out/r02a_rest.log 2>&1; echo "rest rc=$?"; tail -5 gpurun_out/r02a_rest.log
timeout 300 python bench.py > gpurun_out/r02a_bench.json 2> gpurun_out/r02a_bench.err; echo "bench rc=$?"; tail -c 600 gpurun_out/r02a_bench.json
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
