#!/bin/bash
mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_gpu_ops.py -q -k "pack or unpack or merge" > gpurun_out/pack_tests.log 2>&1; echo "pack tests rc=$?"; tail -4 gpurun_out/pack_tests.log
timeout 200 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --per-gpu-batch 8 > gpurun_out/pack_b8.json 2> gpurun_out/pack_b8.err; python -c "
import json; d=json.loads(open('gpurun_out/pack_b8.json').read().strip().splitlines()[-1]); print('batch 8: ms', d['ms_per_step'])"
timeout 200 python tools/bench_extra.py --only pix2pix --steps 10 > gpurun_out/pack_p2p.jsonl 2> gpurun_out/pack_p2p.err; cat gpurun_out/pack_p2p.jsonl
CMD="python bench.py --per-gpu-batch 8 --steps 1 --warmup 3 --no-cpu-baseline --no-graph"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none --profile-from-start off -k regex:"pack|rmsprop" --csv --log-file gpurun_out/launches_pack.csv $CMD > gpurun_out/ncu_pack.log 2>&1
python tools/launch_summary.py gpurun_out/launches_pack.csv
