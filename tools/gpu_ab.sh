#!/bin/bash
# A/B of the conv kernel switches on the bench (ms_per_step + the three microbenchmarked conv passes)
mkdir -p gpurun_out
for cfg in "" "GSB_TC_FULLHALO=0" "GSB_TC_CLUSTER=0" "GSB_TC_BN256=0" "GSB_TC_FULLHALO=0 GSB_TC_CLUSTER=0"; do
  env $cfg python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/ab.log 2>&1
  python - "$cfg" <<'PY'
import json,sys
d=json.loads(open('gpurun_out/ab.log').read().strip().splitlines()[-1])
r=d['roofline']
print('%-40s ms=%.2f fprop64=%.0f %s' % (sys.argv[1] or 'default', d['ms_per_step'], r['achieved'], {k:round(v) for k,v in r['other_kernels_tflops'].items()}))
PY
done
