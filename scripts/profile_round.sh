#!/bin/bash
# Round profile under gpurun (one GPU): launch list of the bench command + ncu --set full captures of the
# hot kernels, each only after the same command has exited 0 without ncu.  Outputs in gpurun_out/.
mkdir -p gpurun_out
TAG=${TAG:-r02}
set -x
python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-extras > gpurun_out/${TAG}_plain_bench.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/${TAG}_launches.csv \
  python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-extras > gpurun_out/${TAG}_ncu_ll.log 2>&1
for op in ${OPS:-attn fc1 tailconv ln}; do
  case $op in
    attn) k=attn_fwd_v2 ;;
    fc1)  k=gemm_tc2 ;;
    tailconv) k=gemm_tc_kernel ;;
    tail_fused) k=depth_tail ;;
    attn_exact) k=attn_fwd_v2 ;;
    ln)   k=layernorm ;;
    *)    k=$op ;;
  esac
  timeout 300 python scripts/bench_ops.py $op > gpurun_out/${TAG}_plain_$op.log 2>&1 &&
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$k -s 3 -c 1 -f -o gpurun_out/${TAG}_$op \
    python scripts/bench_ops.py $op > gpurun_out/${TAG}_ncu_$op.log 2>&1
done
ls -la gpurun_out | tail -20
