#!/bin/bash
# r02x: ncu --set full of every halo-kernel launch of one step (roofline kernel <256,2> and the CTA-pair kernels <64,2,1> / <128,2,1>)
mkdir -p gpurun_out
bash tools/gpu_ncu_full.sh r02x_full_halo "conv_tc_halo_kernel" 34 0
python - <<'PY'
import csv
rows=list(csv.reader(open('gpurun_out/r02x_full_halo.raw.csv')))
hdr=rows[0]; idx={h:i for i,h in enumerate(hdr)}
def g(r,k): 
    return r[idx[k]] if k in idx else ''
for r in rows[2:]:
    print(g(r,'Kernel Name')[:44], g(r,'launch__grid_size'), 'us', g(r,'gpu__time_duration.sum'), 'tensor%', g(r,'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed') or g(r,'sm__inst_executed_pipe_tensor.avg.pct_of_peak_sustained_active'), 'rd', g(r,'dram__bytes_read.sum'), 'wr', g(r,'dram__bytes_write.sum'))
PY
