#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
O=gpurun_out
for op in fc1 fc2 attn; do
  case $op in attn) k=attn_fwd_v2 ;; *) k=gemm_tc2 ;; esac
  timeout 300 python scripts/bench_ops.py $op > $O/c6_plain_$op.log 2>&1 &&
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$k -s 3 -c 1 -f -o $O/c6_$op python scripts/bench_ops.py $op > $O/c6_ncu_$op.log 2>&1
  tail -1 $O/c6_plain_$op.log
done
for rep in 1 2; do timeout -s KILL 400 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > $O/c6_bench_$rep.json 2> $O/c6_bench_$rep.err; python -c "import json; d=json.load(open('$O/c6_bench_$rep.json')); print(round(d['value'],1), d['ms_per_step'], d['clocks']['sm_mhz'], d['breakdown_ms_per_step'], d['roofline']['achieved'])"; done
ls -la $O/*.ncu-rep
