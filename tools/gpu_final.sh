#!/bin/bash
# Round-end GPU call: new tests first (Betty driver over the CUDA modules, unet_128 / n_layers variants, optimisers), then
# the bench line, the driver-based GenSeg iteration, and the rest of the GPU suite with whatever time is left.
mkdir -p gpurun_out
T0=$(date +%s)
el() { echo "[t+$(( $(date +%s) - T0 ))s] $*"; }
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
timeout 300 python -m pytest tests/test_gpu_betty.py -q -s > gpurun_out/q_betty.log 2>&1; el "betty rc=$?"; tail -4 gpurun_out/q_betty.log
timeout 300 python -m pytest tests/test_gpu_pix2pix.py -q -s -k variants > gpurun_out/q_variants.log 2>&1; el "variants rc=$?"; tail -4 gpurun_out/q_variants.log
timeout 200 python -m pytest tests/test_gpu_ops.py -q -k "multi_tensor or rmsprop or adam or optim" > gpurun_out/q_optim.log 2>&1; el "optim rc=$?"; tail -3 gpurun_out/q_optim.log
timeout 420 python bench.py > gpurun_out/q_bench.json 2> gpurun_out/q_bench.err; el "bench rc=$?"; tail -c 600 gpurun_out/q_bench.json; tail -3 gpurun_out/q_bench.err
timeout 240 python tools/bench_extra.py --only betty --steps 5 > gpurun_out/q_betty_bench.jsonl 2> gpurun_out/q_betty_bench.err; el "betty bench rc=$?"; cat gpurun_out/q_betty_bench.jsonl; tail -3 gpurun_out/q_betty_bench.err
timeout 120 python tools/bench_extra.py --only e2e --steps 5 > gpurun_out/q_e2e.jsonl 2> gpurun_out/q_e2e.err; el "e2e rc=$?"; cat gpurun_out/q_e2e.jsonl
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/q_smoke.log 2>&1; el "smoke rc=$?"; tail -2 gpurun_out/q_smoke.log
timeout 600 python -m pytest tests -m gpu -q -x --deselect tests/test_gpu_betty.py --durations=8 > gpurun_out/q_full.log 2>&1; el "full rc=$?"; tail -14 gpurun_out/q_full.log
