#!/bin/bash
# A/B of programmatic dependent launch (GSB_PDL) on the bench step at 64 and 8 images per GPU, then the GPU suite with it on.
mkdir -p gpurun_out
for pdl in 0 1; do
  for b in 64 8; do
    GSB_PDL=$pdl timeout 200 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --per-gpu-batch $b > gpurun_out/pdl_${pdl}_b$b.json 2> gpurun_out/pdl_${pdl}_b$b.err
    python - $pdl $b <<'PY'
import json,sys
try:
    d=json.loads(open('gpurun_out/pdl_%s_b%s.json'%(sys.argv[1],sys.argv[2])).read().strip().splitlines()[-1])
    print('GSB_PDL=%s batch %s: ms_per_step %.3f value %.1f e2e %.1f graph=%s clocks %s' % (sys.argv[1], sys.argv[2], d['ms_per_step'], d['value'], d['e2e']['value'], d['config'].get('cuda_graph'), d['clocks']['sm_mhz']))
except Exception as e:
    print('failed', sys.argv[1:], e); print(open('gpurun_out/pdl_%s_b%s.err'%(sys.argv[1],sys.argv[2])).read()[-1500:])
PY
  done
done
timeout 400 python -m pytest tests -m gpu -q -x > gpurun_out/pdl_tests.log 2>&1; echo "tests rc=$?"; tail -5 gpurun_out/pdl_tests.log
