#!/bin/bash
# r02p: tensor-core kernels of the generator's CW -> 1 transposed 8x8 cell (thin_tc.cu): op test under compute-sanitizer first
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_ops.py -m gpu -q -x -k "thin_tc" > gpurun_out/r02p_thin_tc.log 2>&1; echo "thin_tc rc=$?"; tail -15 gpurun_out/r02p_thin_tc.log | cut -c1-300
GSB_THIN_TC=1 timeout 900 python -m pytest tests/test_gpu_pix2pix.py tests/test_gpu_protocol.py -m gpu -q -x -k "p2p or pix2pix" > gpurun_out/r02p_p2p_tests.log 2>&1; echo "p2p tests (thin_tc on) rc=$?"; tail -4 gpurun_out/r02p_p2p_tests.log | cut -c1-300
for TT in 0 1; do
GSB_THIN_TC=$TT timeout 400 python bench.py --workload pix2pix --steps 10 --warmup 3 --no-cpu-baseline --no-hbm-roofline > gpurun_out/r02p_p2p_tt$TT.json 2> gpurun_out/r02p_p2p_tt$TT.err; echo "p2p thin_tc=$TT rc=$?"; tail -2 gpurun_out/r02p_p2p_tt$TT.err
python -c "
import json; d=json.loads(open('gpurun_out/r02p_p2p_tt$TT.json').read().strip().splitlines()[-1]); print('thin_tc',$TT,'ms',round(d['ms_per_step'],3),'value',round(d['value'],1), d['gpu_launches'])"
done
GSB_THIN_TC=1 bash tools/gpu_list_wl.sh r02p_p2p pix2pix 2>/dev/null | grep -i "thin\|total\|col2im" | head
