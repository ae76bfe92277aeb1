#!/bin/bash
# r02m: BatchNorm + ReLU + MaxPool fused at the encoder seam (forward) and MaxPool backward folded into the BatchNorm backward
# statistics pass: op tests, whole suite, A/B bench, launch list, Pix2Pix launch list
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_ops.py -m gpu -q -x -k "pool" > gpurun_out/r02m_ops.log 2>&1; echo "ops rc=$?"; tail -3 gpurun_out/r02m_ops.log
timeout 2400 python -m pytest tests -m gpu -q -x > gpurun_out/r02m_suite.log 2>&1; echo "suite rc=$?"; tail -4 gpurun_out/r02m_suite.log
for B in 0 8; do for PF in 0 1; do
GSB_POOL_FUSION=$PF timeout 400 python bench.py --no-cpu-baseline --no-hbm-roofline --per-gpu-batch $B > gpurun_out/r02m_bench_pf${PF}_b$B.json 2> gpurun_out/r02m_bench_pf${PF}_b$B.err; echo "bench pool-fusion=$PF b=$B rc=$?"; tail -2 gpurun_out/r02m_bench_pf${PF}_b$B.err
python -c "
import json; d=json.loads(open('gpurun_out/r02m_bench_pf${PF}_b$B.json').read().strip().splitlines()[-1]); print('pool-fusion',$PF,'b',$B,'ms',round(d['ms_per_step'],3),'value',round(d['value'],1), d['clocks']['sm_mhz'], d['gpu_launches'])"
done; done
bash tools/gpu_list.sh r02m_b64 0 | head -34
bash tools/gpu_list_wl.sh r02m_p2p pix2pix | head -50
