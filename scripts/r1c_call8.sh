#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout -s KILL 400 python -m pytest tests/test_gemm_gpu.py -q -m gpu -x 2>&1 | tail -3
timeout -s KILL 300 python scripts/bench_ops.py mm3_geglu --env VDA_GEMM_NO_TMA_EPI=0,4 2>&1 | tail -3
timeout -s KILL 600 python -m pytest tests/test_model_gpu.py -q -m gpu -x 2>&1 | tail -2
