#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
O=gpurun_out
timeout 300 python scripts/bench_ops.py mm3_qkv > $O/c6_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_tc2 -s 3 -c 1 -f -o $O/c6_mm3_qkv python scripts/bench_ops.py mm3_qkv > $O/c6_ncu.log 2>&1
tail -3 $O/c6_ncu.log; ls -la $O/*.ncu-rep
