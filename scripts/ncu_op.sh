#!/bin/bash
# ncu --set full (+ source) of one bench_ops op:  OP=tail_fused K=depth_tail TAG=r02 bash scripts/ncu_op.sh
mkdir -p gpurun_out
TAG=${TAG:-r02}
python scripts/bench_ops.py $OP > gpurun_out/${TAG}_plain_$OP.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:$K -s ${SKIP:-3} -c 1 -f -o gpurun_out/${TAG}_$OP \
    python scripts/bench_ops.py $OP > gpurun_out/${TAG}_ncu_$OP.log 2>&1
tail -2 gpurun_out/${TAG}_plain_$OP.log
ls -la gpurun_out/${TAG}_$OP.ncu-rep
