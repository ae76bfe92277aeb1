#!/bin/bash
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q -x > gpurun_out/r02y_suite.log 2>&1; echo "suite rc=$?"; tail -6 gpurun_out/r02y_suite.log | cut -c1-300
for EF in 0 1; do
GSB_EVAL_FOLD=$EF timeout 400 python bench.py --workload eval --steps 10 --warmup 3 --no-cpu-baseline --no-hbm-roofline > gpurun_out/r02y_eval_fold$EF.jsonl 2> gpurun_out/r02y_eval_fold$EF.err; echo "eval fold=$EF rc=$?"; tail -2 gpurun_out/r02y_eval_fold$EF.err | cut -c1-300
python -c "
import json
for l in open('gpurun_out/r02y_eval_fold$EF.jsonl'):
    if l.startswith('{'):
        d=json.loads(l); print('fold',$EF, d['config']['workload'][-32:], 'ms', round(d['ms_per_step'],3), 'img/s', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'launches', d['gpu_launches'])"
done
