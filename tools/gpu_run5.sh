#!/bin/bash
# r02e (2 GPUs): whole suite incl. the 2-rank test, generator fp32 diag, launch lists, 1- and 2-GPU bench lines of unet / pix2pix
mkdir -p gpurun_out
nvidia-smi -L
timeout 1500 python -m pytest tests -m gpu -q -x --deselect tests/test_gpu_protocol.py -k "not production" > gpurun_out/r02e_suite.log 2>&1; echo "suite rc=$?"; tail -12 gpurun_out/r02e_suite.log
timeout 900 python -m pytest tests/test_gpu_protocol.py -m gpu -q -s > gpurun_out/r02e_protocol.log 2>&1; echo "protocol rc=$?"; tail -6 gpurun_out/r02e_protocol.log
timeout 300 python tools/p2p_diag.py 2>&1 | grep -v initialize | tee gpurun_out/r02e_p2p_diag.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29571 tools/dp_check.py 2>&1 | grep -v "initialize\|Warning\|warn" | tee gpurun_out/r02e_dp_check.log | tail -12
bash tools/gpu_list.sh r02e_b64 0 | head -30
bash tools/gpu_list.sh r02e_b8 8 | head -34
timeout 400 python bench.py > gpurun_out/r02e_bench1.json 2> gpurun_out/r02e_bench1.err; echo "bench1 rc=$?"; tail -3 gpurun_out/r02e_bench1.err
for W in unet pix2pix; do
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29572 bench.py --gpus 2 --workload $W --steps 10 --warmup 3 > gpurun_out/r02e_${W}_2gpu.json 2> gpurun_out/r02e_${W}_2gpu.err; echo "$W 2gpu rc=$?"; tail -2 gpurun_out/r02e_${W}_2gpu.err
done
timeout 300 python bench.py --workload pix2pix --steps 10 > gpurun_out/r02e_pix2pix_1gpu.json 2> gpurun_out/r02e_pix2pix_1gpu.err; echo "p2p1 rc=$?"
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r02e_*gpu.json'))+['gpurun_out/r02e_bench1.json']:
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, 'ms', round(d['ms_per_step'],3), 'value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'launches', d['gpu_launches'], d['clocks'])
    except Exception as e: print(f, 'ERR', e)
PY
