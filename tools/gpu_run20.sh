#!/bin/bash
# r02t: GenSeg iteration as one CUDA graph (A/B against eager)
mkdir -p gpurun_out
for NG in 0 1; do
EXTRA=""; if [ "$NG" = "1" ]; then EXTRA="--no-graph"; fi
timeout 500 python bench.py --workload genseg --steps 10 --warmup 3 --no-cpu-baseline --no-hbm-roofline $EXTRA > gpurun_out/r02t_genseg_nograph$NG.json 2> gpurun_out/r02t_genseg_nograph$NG.err; echo "genseg no-graph=$NG rc=$?"; tail -3 gpurun_out/r02t_genseg_nograph$NG.err | cut -c1-400
python -c "
import json; d=json.loads(open('gpurun_out/r02t_genseg_nograph$NG.json').read().strip().splitlines()[-1]); print('no-graph',$NG,'ms',round(d['ms_per_step'],3),'value',round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'graph', d['config']['cuda_graph'], d['gpu_launches'])"
done
