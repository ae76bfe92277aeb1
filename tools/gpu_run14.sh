#!/bin/bash
# r02n: pool fusion fwd + bwd, parallel planar bias gradient, templated seg_loss kernels: whole suite, bench lines of all workloads
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q -x > gpurun_out/r02n_suite.log 2>&1; echo "suite rc=$?"; tail -4 gpurun_out/r02n_suite.log
timeout 500 python bench.py > gpurun_out/r02n_bench.json 2> gpurun_out/r02n_bench.err; echo "bench rc=$?"; tail -2 gpurun_out/r02n_bench.err
python - <<'PY'
import json
for f in ('r02n_bench',):
    try:
        d=json.loads(open('gpurun_out/%s.json'%f).read().strip().splitlines()[-1])
        print(f,'ms', round(d['ms_per_step'],3), 'value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), d['clocks'], {k: round(v) for k,v in d['roofline']['other_kernels_tflops'].items()}, round(d['roofline']['achieved']), d.get('cpu_baseline'))
        if 'roofline_hbm' in d:
            for k,v in d['roofline_hbm']['kernels'].items(): print('   %-60s %7.1f us frac %.3f' % (k[:60], v['us'], v['frac']))
    except Exception as e: print(f, 'ERR', e)
PY
bash tools/gpu_scale.sh r02n "1" "pix2pix genseg eval"
bash tools/gpu_list_wl.sh r02n_p2p pix2pix | head -24
