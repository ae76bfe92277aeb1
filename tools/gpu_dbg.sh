#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/tests_all.log 2>&1; echo "tests rc=$?"; tail -4 gpurun_out/tests_all.log
timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_b64.log 2>&1; grep '^{"metric"' gpurun_out/bench_b64.log | cut -c1-60,95-200; tail -3 gpurun_out/bench_b64.log | grep -i "error\|Traceback"
python - <<'PY'
import json
for l in open('gpurun_out/bench_b64.log'):
    if l.startswith('{"metric"'):
        d=json.loads(l); print('value', d['value'], 'e2e', d['e2e'], d['clocks'])
PY
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-graph"
timeout 600 $CMD > gpurun_out/plain_x.log 2>&1 &&
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches_x.csv $CMD > gpurun_out/ncu_list_x.log 2>&1
echo "ncu list rc=$?"
