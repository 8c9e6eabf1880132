#!/bin/bash
# 1->N scaling of the headline bench on one box + sharded-driver check.  Usage: scale_run.sh "2 4 8"
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout -s KILL 400 python bench.py --gpus 1 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/scale_1.log 2>&1
tail -n 1 gpurun_out/scale_1.log | cut -c1-160
for n in ${1:-2}; do
  timeout -s KILL 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) \
    bench.py --gpus $n --steps 5 --warmup 3 > gpurun_out/scale_$n.log 2>&1
  echo "N=$n rc=$?"; tail -n 1 gpurun_out/scale_$n.log | cut -c1-160
  timeout -s KILL 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + n)) \
    scripts/mgpu_check.py > gpurun_out/mgpu_$n.log 2>&1
  echo "mgpu N=$n rc=$?"; grep "rank 0" gpurun_out/mgpu_$n.log
done
