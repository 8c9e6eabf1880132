#!/bin/bash
# Staged GPU check run under gpurun: every stage has its own timeout so a hung kernel cannot eat
# the whole lease.  Logs land in gpurun_out/.
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
run() { # name timeout cmd...
  local name=$1 t=$2; shift 2
  echo "=== $name" | tee -a gpurun_out/ci.log
  timeout -s KILL $t "$@" > gpurun_out/$name.log 2>&1
  local rc=$?
  echo "rc=$rc" | tee -a gpurun_out/ci.log
  tail -n ${TAILN:-15} gpurun_out/$name.log | tee -a gpurun_out/ci.log
  return $rc
}
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv | tee -a gpurun_out/ci.log
STAGES=${STAGES:-"gemm kernels model bench"}
for s in $STAGES; do
  case $s in
    gemm)    run gemm 420 python -m pytest tests/test_gemm_gpu.py -q -m gpu -x -s ;;
    kernels) run kernels 420 python -m pytest tests/test_kernels_gpu.py -q -m gpu -x -s ;;
    fp32)    run fp32 900 python -m pytest tests/test_fp32_gpu.py -q -m gpu -x -s ;;
    model)   run model 900 python -m pytest tests/test_model_gpu.py -q -m gpu -s ;;
    all)     run all 1500 python -m pytest tests -q -m gpu -s ;;
    smoke)   run smoke 300 python __graft_entry__.py --smoke ;;
    bench)   run bench 900 python bench.py --steps 5 --warmup 3 ;;
    benchfast) run benchfast 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline ;;
    refarm)  run refarm 600 python bench.py --impl reference --steps 1 --warmup 1 ;;
  esac
done
