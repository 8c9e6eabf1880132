#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 900 python -m pytest tests/test_gemm_gpu.py tests/test_model_gpu.py -q -m gpu -x 2>&1 | tail -4 | tee gpurun_out/c5_tests.log
for h in 0 2 0 2; do
  VDA_CONV_HALO2=$h python bench.py --no-extras --no-cpu-baseline --steps 8 --warmup 3 > gpurun_out/c5_bench_h$h.json 2>gpurun_out/c5_bench_h$h.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/c5_bench_h$h.json").read().strip().split("\n")[-1])
print("HALO2=$h: %.1f fps  conv %.2f ms (%.0f TF/s) gemm %.2f attn %.2f" % (d["value"], d["breakdown_ms_per_step"]["conv3x3"], d["breakdown_tflops"]["conv3x3"], d["breakdown_ms_per_step"]["gemm"], d["breakdown_ms_per_step"]["attn"]))
PY
done 2>&1 | tee gpurun_out/c5_ab_halo2.log
timeout 900 compute-sanitizer --tool memcheck --log-file gpurun_out/r02_sanitizer_memcheck.log python scripts/sanitize_small.py > gpurun_out/r02_sanitizer_memcheck_stdout.log 2>&1; echo "memcheck rc=$?"
tail -5 gpurun_out/r02_sanitizer_memcheck.log; tail -3 gpurun_out/r02_sanitizer_memcheck_stdout.log
