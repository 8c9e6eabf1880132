#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
O=gpurun_out
timeout -s KILL 300 python -m pytest tests/test_kernels_gpu.py -q -m gpu -x -k "bilinear" 2>&1 | tail -3
for rep in 1 2; do
for lib in libvda_b200.so libvda_b200_rows16.so libvda_b200_rows32.so; do
  echo "== $lib"; VDA_LIB_NAME=$lib timeout -s KILL 300 python scripts/bench_ops.py up296 up518 2>&1 | tail -3
done; done | tee $O/c3_bilinear.log
