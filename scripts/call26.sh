#!/bin/bash
# cooperative-box epilogue (bf16 residuals / ReLU'd copy): parity, then A/B against the element-wise path
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
O=gpurun_out
timeout -s KILL 400 python -m pytest tests/test_gemm_gpu.py -x -q -m gpu 2>&1 | tail -6 | tee $O/c26_tests.log
timeout -s KILL 300 python scripts/bench_ops.py conv148 conv148_r2 conv148_r1 conv148_orelu mm3_res1 gemm1024_res1 --env VDA_GEMM_NO_TMA_EPI=5,0 2>&1 | tee $O/c26_ab_coop.log
