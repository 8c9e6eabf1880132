#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
O=gpurun_out
for op in oc1; do
  timeout 300 python scripts/bench_ops.py $op > $O/c7_plain_$op.log 2>&1 &&
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_tc -s 3 -c 1 -f -o $O/c7_$op python scripts/bench_ops.py $op > $O/c7_ncu_$op.log 2>&1
  tail -1 $O/c7_plain_$op.log
done
