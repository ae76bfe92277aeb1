#!/bin/bash
# scaling run on one box: bench.py at N = 8, 4, 2, 1 ranks (+ N-rank == 1-rank parity at the largest N)
mkdir -p gpurun_out
NMAX=${1:-8}
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NMAX --master-addr 127.0.0.1 --master-port 29511 tools/dp_check.py > gpurun_out/dp_check_$NMAX.log 2>&1; echo "dp_check rc=$?"; grep world gpurun_out/dp_check_$NMAX.log
for N in $NMAX 4 2; do
  [ $N -gt $NMAX ] && continue
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2952$N bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/bench_scale$N.log 2>&1; echo "bench$N rc=$?"
  grep '^{"metric"' gpurun_out/bench_scale$N.log | cut -c1-60,95-200
done
timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_scale1.log 2>&1; grep '^{"metric"' gpurun_out/bench_scale1.log | cut -c1-60,95-200
