#!/bin/bash
# One GPU call: the run-vs-run tests touched last + ncu --set full of proj / the two slowest 3x3 conv flavours.
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_model_gpu.py -q -m gpu -k "original_size or pipelined" > gpurun_out/r02f_pytest_subset.log 2>&1
tail -3 gpurun_out/r02f_pytest_subset.log
for OP in ${OPS:-proj conv148_r2 oc1}; do
  OP=$OP K=gemm_tc TAG=r02f bash scripts/ncu_op.sh
done
