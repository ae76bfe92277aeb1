#!/bin/bash
# r02h: CTA-pair halo kernel A/B, fold v3, stem prefetch, head k64, 2x2-block upsample kernels
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_ops.py -m gpu -q -x -k "tile_shapes or upsample" > gpurun_out/r02h_pair.log 2>&1; echo "pair+upsample rc=$?"; tail -4 gpurun_out/r02h_pair.log
timeout 600 python -m pytest tests/test_gpu_ops.py -m gpu -q -x -k "production" > gpurun_out/r02h_prod.log 2>&1; echo "prod rc=$?"; tail -3 gpurun_out/r02h_prod.log
timeout 1500 python -m pytest tests -m gpu -q -x --deselect tests/test_gpu_protocol.py -k "not production and not tile_shapes" > gpurun_out/r02h_suite.log 2>&1; echo "suite rc=$?"; tail -4 gpurun_out/r02h_suite.log
timeout 900 python -m pytest tests/test_gpu_protocol.py -m gpu -q > gpurun_out/r02h_protocol.log 2>&1; echo "protocol rc=$?"; tail -3 gpurun_out/r02h_protocol.log
for PM in 0 1 2; do
GSB_TC_PAIR=$PM timeout 400 python bench.py --no-cpu-baseline > gpurun_out/r02h_bench_pair$PM.json 2> gpurun_out/r02h_bench_pair$PM.err; echo "bench pair=$PM rc=$?"; tail -2 gpurun_out/r02h_bench_pair$PM.err
done
python - <<'PY'
import json
for pm in ('pair0','pair1','pair2'):
    try:
        d=json.loads(open('gpurun_out/r02h_bench_%s.json'%pm).read().strip().splitlines()[-1])
        print(pm,'ms', round(d['ms_per_step'],3), 'value', round(d['value'],1), d['clocks']['sm_mhz'], {k: round(v) for k,v in d['roofline']['other_kernels_tflops'].items()}, round(d['roofline']['achieved']))
        if pm in ('pair1',):
            for k,v in d['roofline_hbm']['kernels'].items(): print('   %-60s %7.1f us frac %.3f' % (k[:60], v['us'], v['frac']))
    except Exception as e: print(pm, 'ERR', e)
PY
for B in 0 8; do for WS in 0 1; do
GSB_WGRAD_STREAM=$WS timeout 400 python bench.py --no-cpu-baseline --no-hbm-roofline --per-gpu-batch $B > gpurun_out/r02h_bench_ws${WS}_b$B.json 2> gpurun_out/r02h_bench_ws${WS}_b$B.err; echo "bench wgrad-stream=$WS b=$B rc=$?"; tail -2 gpurun_out/r02h_bench_ws${WS}_b$B.err
python -c "
import json; d=json.loads(open('gpurun_out/r02h_bench_ws${WS}_b$B.json').read().strip().splitlines()[-1]); print('ws',$WS,'b',$B,'ms',round(d['ms_per_step'],3),'value',round(d['value'],1))"
done; done
GSB_WGRAD_STREAM=1 timeout 600 python -m pytest tests/test_gpu_unet.py tests/test_gpu_protocol.py -m gpu -q -x -k "not p2p and not pix2pix" > gpurun_out/r02h_ws_tests.log 2>&1; echo "ws tests rc=$?"; tail -3 gpurun_out/r02h_ws_tests.log
bash tools/gpu_list.sh r02h_b64 0 | head -32
bash tools/gpu_list.sh r02h_b8 8 | head -32
