#!/bin/bash
# ncu --set full with source-level stall data of the spatial attention kernel (one launch of the vitl block shape)
mkdir -p gpurun_out
TAG=${TAG:-r02}
python scripts/bench_ops.py attn > gpurun_out/${TAG}_plain_attn.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_fwd -s 3 -c 1 -f -o gpurun_out/${TAG}_attn \
    python scripts/bench_ops.py attn > gpurun_out/${TAG}_ncu_attn.log 2>&1
tail -3 gpurun_out/${TAG}_plain_attn.log
ls -la gpurun_out/*.ncu-rep
