#!/bin/bash
# r02k: one-wave weight-gradient splits A/B (64 and 8 images per GPU), determinism + production tests with it, ncu --set full
# of the bandwidth kernels still under 0.7 of the HBM roofline
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ops.py tests/test_gpu_unet.py -m gpu -q -x -k "production or deterministic or reproducible or wgrad" > gpurun_out/r02k_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/r02k_tests.log
for B in 0 8; do for OW in 0 1; do
GSB_WGRAD_ONE_WAVE=$OW timeout 400 python bench.py --no-cpu-baseline --no-hbm-roofline --per-gpu-batch $B > gpurun_out/r02k_bench_ow${OW}_b$B.json 2> gpurun_out/r02k_bench_ow${OW}_b$B.err; echo "bench one-wave=$OW b=$B rc=$?"; tail -2 gpurun_out/r02k_bench_ow${OW}_b$B.err
python -c "
import json; d=json.loads(open('gpurun_out/r02k_bench_ow${OW}_b$B.json').read().strip().splitlines()[-1]); print('one-wave',$OW,'b',$B,'ms',round(d['ms_per_step'],3),'value',round(d['value'],1), d['clocks']['sm_mhz'], {k: round(v) for k,v in d['roofline']['other_kernels_tflops'].items()})"
done; done
bash tools/gpu_list.sh r02k_b8 8 | head -16
bash tools/gpu_ncu_full.sh r02k_full_bw "stem_tc_fwd|upsample2_bwd_sep|upsample2_fwd_blk|seg_loss_reduce_kernel|fold_slices|head_bwd" 14 0
