#!/bin/bash
# A/B of the tail chunking modes (engine.head): headline bench without extras, interleaved.
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
for rep in 1 2; do
for cfg in "0 1" "1 1" "1 2" "2 1" "2 2"; do
  set -- $cfg
  VDA_TAIL_MODE=$1 VDA_TAIL_CHUNK=$2 python bench.py --no-extras --no-cpu-baseline --steps 8 --warmup 3 > gpurun_out/ab_tail_$1_$2_$rep.json 2> gpurun_out/ab_tail_$1_$2_$rep.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/ab_tail_$1_$2_$rep.json").read().strip().split("\n")[-1])
print("mode $1 chunk $2 rep $rep: %.1f fps  conv %.2f ms bilinear %.2f ms" % (d["value"], d["breakdown_ms_per_step"]["conv3x3"], d["hbm_kernels"]["bilinear"]["ms_per_step"]))
PY
done
done 2>&1 | tee gpurun_out/ab_tail.log
VDA_TAIL_MODE=2 VDA_TAIL_CHUNK=1 python -m pytest tests/test_model_gpu.py -q -m gpu -x 2>&1 | tail -3 | tee -a gpurun_out/ab_tail.log
VDA_TAIL_MODE=1 VDA_TAIL_CHUNK=2 python -m pytest tests/test_model_gpu.py -q -m gpu -x 2>&1 | tail -3 | tee -a gpurun_out/ab_tail.log
