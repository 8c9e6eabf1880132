#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
O=gpurun_out
( time timeout -s KILL 1500 python -m pytest tests/ -x -q -m gpu ) > $O/c27_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 $O/c27_pytest_gpu.log
for v in 5 0 5 0; do VDA_GEMM_NO_TMA_EPI=$v timeout -s KILL 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-extras 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().split('\n')[-1]); print('NO_TMA_EPI=$v', round(d['value'],1), 'fps', d['breakdown_ms_per_step'])" | tee -a $O/c27_ab_window.log; done
