#!/bin/bash
mkdir -p gpurun_out
for a in "" "--no-graph" "--per-gpu-batch 8" "--per-gpu-batch 8 --no-graph"; do
timeout 600 python bench.py $a --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_g.log 2>gpurun_out/bench_g.err; echo "[$a] rc=$?"; tail -2 gpurun_out/bench_g.err | cut -c1-300
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_g.log').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], 'launches', d['gpu_launches'], d['config'].get('cuda_graph'))
PY
done
