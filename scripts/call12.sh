#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 300 python -m pytest tests/test_kernels_gpu.py -q -m gpu -x -k "depth_tail" 2>&1 | tail -4 | tee gpurun_out/c12_test.log
timeout 200 python scripts/bench_ops.py tail_fused tail_unfused 2>&1 | tee gpurun_out/c12_tail_ops.log
for f in 0 1 0 1; do
  VDA_TAIL_FUSED=$f python bench.py --no-extras --no-cpu-baseline --steps 8 --warmup 3 > gpurun_out/c12_bench_f$f.json 2>gpurun_out/c12_bench_f$f.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/c12_bench_f$f.json").read().strip().split("\n")[-1])
print("TAIL_FUSED=$f: %.1f fps  conv %.2f ms bilinear %.2f ms gemm %.2f attn %.2f" % (d["value"], d["breakdown_ms_per_step"]["conv3x3"], d["hbm_kernels"]["bilinear"]["ms_per_step"], d["breakdown_ms_per_step"]["gemm"], d["breakdown_ms_per_step"]["attn"]))
PY
done 2>&1 | tee gpurun_out/c12_ab_tail_fused.log
OP=fc1 K=gemm_tc2 TAG=r02 bash scripts/ncu_op.sh
