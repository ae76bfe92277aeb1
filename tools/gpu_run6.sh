#!/bin/bash
# r02f: tensor-core stem, fold kernel fix, fp64-accumulated SIMT convolutions
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_ops.py -m gpu -q -x -k "stem or batched_pack or determin or instance" -s > gpurun_out/r02f_new.log 2>&1; echo "new rc=$?"; tail -12 gpurun_out/r02f_new.log
timeout 1500 python -m pytest tests -m gpu -q -x --deselect tests/test_gpu_protocol.py -k "not production" > gpurun_out/r02f_suite.log 2>&1; echo "suite rc=$?"; tail -8 gpurun_out/r02f_suite.log
timeout 900 python -m pytest tests/test_gpu_protocol.py -m gpu -q -s > gpurun_out/r02f_protocol.log 2>&1; echo "protocol rc=$?"; tail -6 gpurun_out/r02f_protocol.log
grep -n "trajectory\|median" gpurun_out/r02f_protocol.log | cut -c1-300
timeout 300 python tools/p2p_diag.py 2>&1 | grep -v initialize | head -4
bash tools/gpu_list.sh r02f_b64 0 | head -34
bash tools/gpu_list.sh r02f_b8 8 | head -34
timeout 400 python bench.py > gpurun_out/r02f_bench.json 2> gpurun_out/r02f_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/r02f_bench.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02f_bench.json').read().strip().splitlines()[-1]); print('ms', round(d['ms_per_step'],3), 'value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'launches', d['gpu_launches'], d['clocks'])
for k,v in d['roofline_hbm']['kernels'].items(): print('   %-60s %7.1f us frac %.3f' % (k[:60], v['us'], v['frac']))
PY
