#!/bin/bash
mkdir -p gpurun_out
CMD="python tools/bench_extra.py --only pix2pix --steps 1"
timeout 600 $CMD > gpurun_out/plain_p2p.log 2>&1 &&
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none --profile-from-start off --csv --log-file gpurun_out/launches_p2p.csv $CMD > gpurun_out/ncu_list_p2p.log 2>&1
echo rc=$?
