#!/bin/bash
# usage: bash tools/gpu_scale.sh TAG "N1 N2 ..." "workloads" : bench lines of every workload at every rank count of ONE box
# (run under gpurun --gpus max(N)); writes gpurun_out/${TAG}_scale.jsonl (one bench JSON line per run, with clocks)
TAG=$1; NS=${2:-"1 2"}; WLS=${3:-"unet pix2pix genseg"}
mkdir -p gpurun_out
nvidia-smi -L | head -8
: > gpurun_out/${TAG}_scale.jsonl
PORT=29600
for W in $WLS; do for N in $NS; do
  PORT=$((PORT+1))
  if [ "$N" = "1" ]; then
    timeout 500 python bench.py --workload $W --steps 10 --warmup 3 --no-cpu-baseline --no-hbm-roofline > gpurun_out/${TAG}_${W}_${N}.json 2> gpurun_out/${TAG}_${W}_${N}.err
  else
    timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $N --workload $W --steps 10 --warmup 3 --no-cpu-baseline --no-hbm-roofline > gpurun_out/${TAG}_${W}_${N}.json 2> gpurun_out/${TAG}_${W}_${N}.err
  fi
  echo "$W N=$N rc=$?"; grep -h '"metric"' gpurun_out/${TAG}_${W}_${N}.json | tail -1 >> gpurun_out/${TAG}_scale.jsonl
done; done
python - <<PY
import json
base = {}
for l in open('gpurun_out/${TAG}_scale.jsonl'):
    d = json.loads(l); w = d['config'].get('workload','')[:28]; n = d['n_gpus']
    base.setdefault(w, d['value'] if n == 1 else None)
    b = base.get(w)
    print('%-30s N=%d  %8.3f ms  %10.1f %s  e2e %10.1f  x%.2f  clocks %s %s' % (w, n, d['ms_per_step'], d['value'], d['unit'][:12], d['e2e']['value'], (d['value']/b if b else float('nan')), d['clocks'].get('sm_mhz'), d['clocks'].get('reasons')))
PY
