#!/bin/bash
# 8-GPU run of BASELINE configs[2]: 1024-frame vitl video, windows sharded over the GPUs; every rank returning the full
# result vs rank 0 only.
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
O=gpurun_out
timeout -s KILL 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29708 scripts/video_bench.py --frames 1024 --reps 3 --result-rank both > $O/g8_video.log 2>&1; echo "video rc=$?"; grep '"workload"' $O/g8_video.log | cut -c1-360
