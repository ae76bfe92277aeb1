#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_protocol.py -m gpu -q -s -k "pix2pix_every" > gpurun_out/r02d_protocol.log 2>&1; echo "protocol rc=$?"; tail -4 gpurun_out/r02d_protocol.log
timeout 600 python -m pytest tests/test_gpu_ops.py tests/test_gpu_pix2pix.py -m gpu -q -x -k "not production" > gpurun_out/r02d_ops.log 2>&1; echo "ops rc=$?"; tail -4 gpurun_out/r02d_ops.log
bash tools/gpu_list.sh r02d_b64 0
bash tools/gpu_list.sh r02d_b8 8
timeout 400 python bench.py > gpurun_out/r02d_bench.json 2> gpurun_out/r02d_bench.err; echo "bench rc=$?"; tail -c 3000 gpurun_out/r02d_bench.json; tail -3 gpurun_out/r02d_bench.err
timeout 300 python bench.py --workload pix2pix --steps 10 > gpurun_out/r02d_pix2pix.json 2> gpurun_out/r02d_pix2pix.err; echo "p2p rc=$?"; tail -c 600 gpurun_out/r02d_pix2pix.json; tail -3 gpurun_out/r02d_pix2pix.err
timeout 300 python bench.py --workload eval --steps 10 > gpurun_out/r02d_eval.json 2> gpurun_out/r02d_eval.err; echo "eval rc=$?"; cut -c1-200 gpurun_out/r02d_eval.json; tail -3 gpurun_out/r02d_eval.err
timeout 300 python bench.py --workload genseg --steps 5 > gpurun_out/r02d_genseg.json 2> gpurun_out/r02d_genseg.err; echo "genseg rc=$?"; cut -c1-300 gpurun_out/r02d_genseg.json; tail -3 gpurun_out/r02d_genseg.err
