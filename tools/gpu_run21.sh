#!/bin/bash
# r02u (2 GPUs): graph replay of the data-parallel Pix2Pix / GenSeg steps (NCCL collectives inside the captured graph)
mkdir -p gpurun_out
for W in pix2pix genseg; do for NG in 0 1; do
EXTRA=""; if [ "$NG" = "1" ]; then EXTRA="--no-graph"; fi
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2961$NG bench.py --gpus 2 --workload $W --steps 10 --warmup 3 --no-cpu-baseline --no-hbm-roofline $EXTRA > gpurun_out/r02u_${W}_2_nograph$NG.json 2> gpurun_out/r02u_${W}_2_nograph$NG.err; echo "$W 2gpu no-graph=$NG rc=$?"; grep -i "capture failed\|error" gpurun_out/r02u_${W}_2_nograph$NG.err | head -3 | cut -c1-300
python -c "
import json; d=json.loads([l for l in open('gpurun_out/r02u_${W}_2_nograph$NG.json') if l.startswith('{')][-1]); print('$W','no-graph',$NG,'ms',round(d['ms_per_step'],3),'value',round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'graph', d['config']['cuda_graph'])"
done; done
