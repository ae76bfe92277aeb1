#!/bin/bash
# final-state validation (pass a tag: tools/gpu_validate.sh r03a): build check, smoke(), whole GPU suite, default bench line, reference arm, launch lists
TAG=${1:-val}
mkdir -p gpurun_out
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/${TAG}_smoke.log | cut -c1-300
timeout 2400 python -m pytest tests -m gpu -q -x > gpurun_out/${TAG}_suite.log 2>&1; echo "suite rc=$?"; tail -4 gpurun_out/${TAG}_suite.log | cut -c1-300
timeout 600 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"; tail -2 gpurun_out/${TAG}_bench.err | cut -c1-300
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${TAG}_bench_reference.json 2> gpurun_out/${TAG}_bench_reference.err; echo "reference arm rc=$?"; cat gpurun_out/${TAG}_bench_reference.json | cut -c1-600
python - <<PY
import json
d=json.loads(open('gpurun_out/${TAG}_bench.json').read().strip().splitlines()[-1])
print('bench ms', round(d['ms_per_step'],3), 'value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), d['clocks'], 'launches', d['gpu_launches'], 'roofline', round(d['roofline']['achieved']), round(d['roofline']['frac'],3), d['cpu_baseline'])
for k,v in d['roofline_hbm']['kernels'].items(): print('   %-60s %7.1f us frac %.3f' % (k[:60], v['us'], v['frac']))
PY
bash tools/gpu_list.sh ${TAG}_b64 0 | head -36
bash tools/gpu_list.sh ${TAG}_b8 8 | head -36
