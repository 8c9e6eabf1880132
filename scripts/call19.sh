#!/bin/bash
# attention v3 (persistent ping-pong): parity tests, then A/B against v2
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
O=gpurun_out
timeout -s KILL 240 python -m pytest tests/test_kernels_gpu.py -x -q -m gpu -k "flash_attention" 2>&1 | tail -15 | tee $O/c19_attn_tests.log
timeout -s KILL 120 python scripts/bench_ops.py attn_exact --env VDA_ATTN_V=2,3 2>&1 | tee $O/c19_attn_ab.log
timeout -s KILL 120 python scripts/bench_ops.py attn_exact --env VDA_ATTN_PP=0,1 2>&1 | tee -a $O/c19_attn_ab.log
timeout -s KILL 120 python scripts/bench_ops.py attn_exact --env VDA_ATTN_POLY=0,4,8 2>&1 | tee -a $O/c19_attn_ab.log
