#!/bin/bash
mkdir -p gpurun_out
timeout 500 python -m pytest tests -m gpu -q -x > gpurun_out/f_full.log 2>&1; echo "full rc=$?"; tail -4 gpurun_out/f_full.log
timeout 300 python bench.py > gpurun_out/f_bench.json 2> gpurun_out/f_bench.err; echo "bench rc=$?"; python -c "
import json; d=json.loads(open('gpurun_out/f_bench.json').read().strip().splitlines()[-1]); print('bench: ms', d['ms_per_step'], 'value', d['value'], 'e2e', d['e2e']['value'], 'clk', d['clocks'])"
timeout 200 python tools/bench_extra.py --only pix2pix --steps 10 2>/dev/null | tail -1 | cut -c1-200
timeout 200 python tools/bench_extra.py --only e2e --steps 5 2>/dev/null | tail -1 | cut -c1-260
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
