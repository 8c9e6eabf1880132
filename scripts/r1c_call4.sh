#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
O=gpurun_out
timeout -s KILL 400 python -m pytest tests/test_gemm_gpu.py -q -m gpu -x 2>&1 | tail -4
timeout -s KILL 300 python scripts/bench_ops.py ct0 ct1 --env VDA_GEMM_NO_TMA_EPI=0,3 2>&1 | tail -6 | tee $O/c4_convt.log
timeout -s KILL 300 python -m pytest tests/test_kernels_gpu.py -q -m gpu -x 2>&1 | tail -3
timeout -s KILL 600 python -m pytest tests/test_model_gpu.py -q -m gpu -x 2>&1 | tail -3
for rep in 1 2; do timeout -s KILL 400 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > $O/c4_bench_$rep.json 2> $O/c4_bench_$rep.err; python -c "import json; d=json.load(open('$O/c4_bench_$rep.json')); print(round(d['value'],1), d['ms_per_step'], d['clocks']['sm_mhz'], d['breakdown_ms_per_step'], {k:round(v['ms_per_step'],2) for k,v in d['hbm_kernels'].items()})"; done
timeout -s KILL 300 python scripts/stream_bench.py --arch vits | cut -c1-200
