#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 900 python -m pytest tests/test_gemm_gpu.py tests/test_model_gpu.py tests/test_dropin_gpu.py -q -m gpu -x 2>&1 | tail -4 | tee gpurun_out/c7_tests.log
for k in 0 1 0 1; do
  if [ $k = 0 ]; then export VDA_GEMM_KSPLIT=0; else unset VDA_GEMM_KSPLIT; fi
  python scripts/stream_bench.py --arch vitl --frames 120 | cut -c1-200
  python scripts/stream_bench.py --arch vits --frames 120 | cut -c1-200
done 2>&1 | tee gpurun_out/c7_stream_ab.log
unset VDA_GEMM_KSPLIT
python bench.py --no-extras --no-cpu-baseline --steps 8 --warmup 3 > gpurun_out/c7_bench.json 2>gpurun_out/c7_bench.err
python - <<PY
import json
d=json.loads(open("gpurun_out/c7_bench.json").read().strip().split("\n")[-1])
print("bench: %.1f fps  conv %.2f gemm %.2f attn %.2f" % (d["value"], d["breakdown_ms_per_step"]["conv3x3"], d["breakdown_ms_per_step"]["gemm"], d["breakdown_ms_per_step"]["attn"]))
PY
