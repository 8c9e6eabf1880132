#!/bin/bash
# no-alloc entry point: alone in a fresh process, then at the end of the whole GPU suite; then the round-end sequence
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
O=gpurun_out
( timeout -s KILL 300 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu ) > $O/c15_noalloc_alone.log 2>&1; echo "alone rc=$?"; tail -3 $O/c15_noalloc_alone.log
( time timeout -s KILL 1500 python -m pytest tests/ -x -q -m gpu ) > $O/c15_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 $O/c15_pytest_gpu.log
( time timeout -s KILL 300 python -c "import __graft_entry__ as g; g.smoke()" ) > $O/c15_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 $O/c15_smoke.log
( time timeout -s KILL 900 python bench.py ) > $O/c15_bench.json 2> $O/c15_bench.err; echo "bench rc=$?"; tail -2 $O/c15_bench.err; cut -c1-600 $O/c15_bench.json
