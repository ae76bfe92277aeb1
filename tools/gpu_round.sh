#!/bin/bash
# One GPU-box round: ops tests, fp32 U-Net parity, tcgen05 diagnostics, bf16 U-Net parity, smoke, short bench.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
timeout 600 python -m pytest tests/test_gpu_ops.py -q -k "not tcgen05" > gpurun_out/ops.log 2>&1; echo "ops rc=$?" 
tail -5 gpurun_out/ops.log
timeout 900 python -m pytest tests/test_gpu_unet.py -q -k "not bf16" > gpurun_out/unet_fp32.log 2>&1; echo "unet_fp32 rc=$?"
tail -5 gpurun_out/unet_fp32.log
timeout 900 python tools/tc_diag.py > gpurun_out/tc_diag.log 2>&1; echo "tc_diag rc=$?"
cat gpurun_out/tc_diag.log
timeout 600 python -m pytest tests/test_gpu_unet.py -q -s -k "bf16" > gpurun_out/unet_bf16.log 2>&1; echo "unet_bf16 rc=$?"
tail -8 gpurun_out/unet_bf16.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.log 2>&1; echo "bench rc=$?"; tail -3 gpurun_out/bench.log
