#!/bin/bash
# attention v3 chained mode (VDA_ATTN_V=4): parity tests, then A/B against v2 over items per CTA
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
O=gpurun_out
VDA_ATTN_V=4 timeout -s KILL 240 python -m pytest tests/test_kernels_gpu.py -x -q -m gpu -k "flash_attention" 2>&1 | tail -5 | tee $O/c22_attn_tests.log
VDA_ATTN_V=4 VDA_ATTN_IPC=3 timeout -s KILL 240 python -m pytest tests/test_kernels_gpu.py -x -q -m gpu -k "flash_attention" 2>&1 | tail -5 | tee -a $O/c22_attn_tests.log
timeout -s KILL 120 python scripts/bench_ops.py attn_exact --env VDA_ATTN_V=2,4 2>&1 | tee $O/c22_attn_ab.log
VDA_ATTN_V=4 timeout -s KILL 120 python scripts/bench_ops.py attn_exact --env VDA_ATTN_IPC=1,2,3,4,6 2>&1 | tee -a $O/c22_attn_ab.log
