#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
O=gpurun_out
timeout -s KILL 300 python -m pytest tests/test_kernels_gpu.py -x -q -m gpu -k "temporal_attn_stream" 2>&1 | tail -8 | tee $O/c29_tests.log
VDA_TSTREAM=1 timeout -s KILL 300 python -m pytest tests/test_kernels_gpu.py -x -q -m gpu -k "temporal_attn_stream" 2>&1 | tail -3 | tee -a $O/c29_tests.log
timeout -s KILL 600 python -m pytest tests/ -x -q -m gpu -k "stream" 2>&1 | tail -4 | tee -a $O/c29_tests.log
for v in 1 0; do for a in vitl vits; do VDA_TSTREAM=$v timeout 300 python scripts/stream_bench.py --arch $a --frames 120 | cut -c1-150 | sed "s/^/TSTREAM=$v /" | tee -a $O/c29_stream.log; done; done
