#!/bin/bash
# r02s: capturable Pix2Pix step (device-side Adam step count + dropout call count), graph replay in bench --workload pix2pix
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_pix2pix.py -m gpu -q -x -k "cuda_graph" > gpurun_out/r02s_graph_test.log 2>&1; echo "graph test rc=$?"; tail -12 gpurun_out/r02s_graph_test.log | cut -c1-400
for NG in 0 1; do
EXTRA=""; if [ "$NG" = "1" ]; then EXTRA="--no-graph"; fi
timeout 400 python bench.py --workload pix2pix --steps 10 --warmup 3 --no-cpu-baseline --no-hbm-roofline $EXTRA > gpurun_out/r02s_p2p_nograph$NG.json 2> gpurun_out/r02s_p2p_nograph$NG.err; echo "p2p no-graph=$NG rc=$?"; tail -3 gpurun_out/r02s_p2p_nograph$NG.err | cut -c1-300
python -c "
import json; d=json.loads(open('gpurun_out/r02s_p2p_nograph$NG.json').read().strip().splitlines()[-1]); print('no-graph',$NG,'ms',round(d['ms_per_step'],3),'value',round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'graph', d['config']['cuda_graph'], d['gpu_launches'])"
done
timeout 1500 python -m pytest tests/test_gpu_pix2pix.py tests/test_gpu_protocol.py tests/test_zz_no_alloc.py -m gpu -q -x > gpurun_out/r02s_p2p_suite.log 2>&1; echo "p2p suite rc=$?"; tail -3 gpurun_out/r02s_p2p_suite.log | cut -c1-300
