#!/bin/bash
# Round-end check as the driver runs it: full GPU test suite, smoke(), default bench (with CPU baseline), reference arm.
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
O=gpurun_out
( time timeout -s KILL 1500 python -m pytest tests/ -x -q -m gpu ) > $O/f_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 $O/f_pytest_gpu.log
( time timeout -s KILL 300 python -c "import __graft_entry__ as g; g.smoke()" ) > $O/f_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 $O/f_smoke.log
( time timeout -s KILL 900 python bench.py ) > $O/f_bench.json 2> $O/f_bench.err; echo "bench rc=$?"; tail -2 $O/f_bench.err; cut -c1-400 $O/f_bench.json
( time timeout -s KILL 600 python bench.py --impl reference --steps 2 --warmup 1 ) > $O/f_bench_ref.json 2> $O/f_bench_ref.err; echo "ref rc=$?"; cut -c1-300 $O/f_bench_ref.json; tail -3 $O/f_bench_ref.err
