#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
O=gpurun_out
for rep in 1 2; do
for lib in libvda_b200_rel.so libvda_b200.so; do
  echo "== $lib"; VDA_LIB_NAME=$lib timeout -s KILL 300 python scripts/bench_ops.py qkv proj fc1 fc2 ct0 ct1 mm3_qkv mm3_out conv148 oc1 2>&1 | tail -11
done; done | tee $O/c5_arrive.log
timeout -s KILL 400 python -m pytest tests/test_gemm_gpu.py -q -m gpu -x 2>&1 | tail -2
timeout -s KILL 600 python -m pytest tests/test_model_gpu.py -q -m gpu -x 2>&1 | tail -2
