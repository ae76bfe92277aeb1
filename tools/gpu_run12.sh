#!/bin/bash
# r02l: one-wave weight-gradient splits, fold kernel v4, TMA-store stem epilogue, upsample tweaks: whole suite, bench lines, lists
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q -x > gpurun_out/r02l_suite.log 2>&1; echo "suite rc=$?"; tail -4 gpurun_out/r02l_suite.log
timeout 500 python bench.py > gpurun_out/r02l_bench.json 2> gpurun_out/r02l_bench.err; echo "bench rc=$?"; tail -2 gpurun_out/r02l_bench.err
timeout 400 python bench.py --no-cpu-baseline --no-hbm-roofline --per-gpu-batch 8 > gpurun_out/r02l_bench_b8.json 2> gpurun_out/r02l_bench_b8.err; echo "bench b8 rc=$?"
python - <<'PY'
import json
for f in ('r02l_bench','r02l_bench_b8'):
    try:
        d=json.loads(open('gpurun_out/%s.json'%f).read().strip().splitlines()[-1])
        print(f,'ms', round(d['ms_per_step'],3), 'value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), d['clocks'], {k: round(v) for k,v in d['roofline']['other_kernels_tflops'].items()}, round(d['roofline']['achieved']), d.get('cpu_baseline'))
        if 'roofline_hbm' in d:
            for k,v in d['roofline_hbm']['kernels'].items(): print('   %-60s %7.1f us frac %.3f' % (k[:60], v['us'], v['frac']))
    except Exception as e: print(f, 'ERR', e)
PY
bash tools/gpu_list.sh r02l_b64 0 | head -34
bash tools/gpu_list.sh r02l_b8 8 | head -34
