#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 600 python -m pytest tests/test_dropin_gpu.py tests/test_model_gpu.py tests/test_fp32_gpu.py -q -m gpu -x -k "stream" 2>&1 | tail -4 | tee gpurun_out/c14_test.log
python scripts/stream_bench.py --arch vitl --frames 120 | cut -c1-160 | tee gpurun_out/c14_stream.log
python scripts/stream_bench.py --arch vits --frames 120 | cut -c1-160 | tee -a gpurun_out/c14_stream.log
timeout 300 python scripts/bench_ops.py proj proj_wf32 proj_bias fc2 --env VDA_GEMM_RED_AS_STORE=0,1 2>&1 | tee gpurun_out/c14_red_vs_store.log
