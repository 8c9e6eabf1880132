#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 400 python -m pytest tests/test_kernels_gpu.py -q -m gpu -x -k "flash_attention" 2>&1 | tail -15 | tee gpurun_out/c9_attn_test.log
timeout 200 python scripts/bench_ops.py attn attn_exact 2>&1 | tee gpurun_out/c9_attn_ops.log
timeout 200 python scripts/bench_ops.py attn --env VDA_ATTN_POLY=0,8 2>&1 | tee -a gpurun_out/c9_attn_ops.log
timeout 600 python -m pytest tests/test_model_gpu.py -q -m gpu -x 2>&1 | tail -3 | tee -a gpurun_out/c9_attn_test.log
for f in 0 1 0 1; do
  VDA_ATTN_OPT=$f python bench.py --no-extras --no-cpu-baseline --steps 8 --warmup 3 > gpurun_out/c9_bench_o$f.json 2>gpurun_out/c9_bench_o$f.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/c9_bench_o$f.json").read().strip().split("\n")[-1])
print("ATTN_OPT=$f: %.1f fps  attn %.2f ms (%.0f TF/s) gemm %.2f conv %.2f" % (d["value"], d["breakdown_ms_per_step"]["attn"], d["breakdown_tflops"]["attn"], d["breakdown_ms_per_step"]["gemm"], d["breakdown_ms_per_step"]["conv3x3"]))
PY
done 2>&1 | tee gpurun_out/c9_ab_attn_opt.log
