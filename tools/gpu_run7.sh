#!/bin/bash
# r02g: CTA-pair (cta_group::2) halo kernel, fold kernel v3, stem prefetch, head_fwd k64
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_ops.py -m gpu -q -x -k "tile_shapes" > gpurun_out/r02g_pair.log 2>&1; echo "pair rc=$?"; tail -8 gpurun_out/r02g_pair.log
timeout 600 python -m pytest tests/test_gpu_ops.py -m gpu -q -x -k "production" > gpurun_out/r02g_prod.log 2>&1; echo "prod rc=$?"; tail -4 gpurun_out/r02g_prod.log
timeout 1500 python -m pytest tests -m gpu -q -x --deselect tests/test_gpu_protocol.py -k "not production and not tile_shapes" > gpurun_out/r02g_suite.log 2>&1; echo "suite rc=$?"; tail -8 gpurun_out/r02g_suite.log
timeout 900 python -m pytest tests/test_gpu_protocol.py -m gpu -q > gpurun_out/r02g_protocol.log 2>&1; echo "protocol rc=$?"; tail -4 gpurun_out/r02g_protocol.log
for PM in 0 1 2; do
GSB_TC_PAIR=$PM timeout 400 python bench.py --no-cpu-baseline > gpurun_out/r02g_bench_pair$PM.json 2> gpurun_out/r02g_bench_pair$PM.err; echo "bench pair=$PM rc=$?"; tail -2 gpurun_out/r02g_bench_pair$PM.err
done
python - <<'PY'
import json
for pm in (0,1,2):
    try:
        d=json.loads(open('gpurun_out/r02g_bench_pair%d.json'%pm).read().strip().splitlines()[-1])
        print('pair',pm,'ms', round(d['ms_per_step'],3), 'value', round(d['value'],1), d['clocks']['sm_mhz'], {k: round(v) for k,v in d['roofline']['other_kernels_tflops'].items()}, round(d['roofline']['achieved']))
        if pm==1:
            for k,v in d['roofline_hbm']['kernels'].items(): print('   %-60s %7.1f us frac %.3f' % (k[:60], v['us'], v['frac']))
    except Exception as e: print(pm, 'ERR', e)
PY
bash tools/gpu_list.sh r02g_b64 0 | head -30
bash tools/gpu_list.sh r02g_b8 8 | head -30
