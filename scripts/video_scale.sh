#!/bin/bash
# BASELINE configs[2]: 1024-frame vitl video through infer_video_depth, windows sharded over N GPUs.
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
for n in ${1:-8}; do
  timeout -s KILL 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29700 + n)) \
    scripts/video_bench.py --frames 1024 --reps 2 > gpurun_out/video_$n.log 2>&1
  echo "N=$n rc=$?"; grep '"workload"' gpurun_out/video_$n.log | tail -n 1
done
