#!/bin/bash
# r02r: thin-convolution grid heuristics (PatchGAN head), programmatic dependent launch A/B on the eager Pix2Pix step
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ops.py tests/test_gpu_pix2pix.py -m gpu -q -x -k "thin or pix2pix or p2p or discriminator" > gpurun_out/r02r_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/r02r_tests.log | cut -c1-300
for PDL in 0 1; do
GSB_PDL=$PDL timeout 400 python bench.py --workload pix2pix --steps 10 --warmup 3 --no-cpu-baseline --no-hbm-roofline > gpurun_out/r02r_p2p_pdl$PDL.json 2> gpurun_out/r02r_p2p_pdl$PDL.err; echo "p2p pdl=$PDL rc=$?"; tail -2 gpurun_out/r02r_p2p_pdl$PDL.err
python -c "
import json; d=json.loads(open('gpurun_out/r02r_p2p_pdl$PDL.json').read().strip().splitlines()[-1]); print('pdl',$PDL,'ms',round(d['ms_per_step'],3),'value',round(d['value'],1), 'e2e', round(d['e2e']['value'],1))"
done
bash tools/gpu_list_wl.sh r02r_p2p pix2pix 2>/dev/null | grep -i "thin\|total" | head
