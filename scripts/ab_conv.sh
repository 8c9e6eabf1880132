#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 600 python -m pytest tests/test_gemm_gpu.py -q -m gpu -x -k "conv3x3" 2>&1 | tail -5 | tee gpurun_out/ab_conv.log
timeout 300 python scripts/bench_ops.py oc1 conv148 --env VDA_CONV_HALO2=0,1,2 2>&1 | tee -a gpurun_out/ab_conv.log
