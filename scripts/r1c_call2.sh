#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
O=gpurun_out
timeout -s KILL 300 python scripts/attn_wave_scan.py > $O/c2_attn_wave.log 2>&1; cat $O/c2_attn_wave.log
for rep in 1 2; do
 for cfg in "VDA_PDL=0 VDA_ENC_CHUNKS=1" "VDA_PDL=1 VDA_ENC_CHUNKS=1" "VDA_PDL=1 VDA_ENC_CHUNKS=2" "VDA_PDL=1 VDA_ENC_CHUNKS=4" "VDA_PDL=0 VDA_ENC_CHUNKS=4"; do
  tag=$(echo $cfg | tr ' =' '__')
  env $cfg timeout -s KILL 400 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > $O/c2_bench_${tag}_$rep.json 2> $O/c2_bench_${tag}_$rep.err
  echo "$cfg rep$rep: $(python -c "import json,sys; d=json.load(open('$O/c2_bench_${tag}_$rep.json')); print(round(d['value'],1), d['ms_per_step'], d['clocks'], d['breakdown_ms_per_step'], {k:round(v['ms_per_step'],2) for k,v in d['hbm_kernels'].items()})")"
 done
done
VDA_ENC_CHUNKS=4 timeout -s KILL 600 python -m pytest tests/test_model_gpu.py -q -m gpu -x > $O/c2_model_chunks4.log 2>&1; tail -3 $O/c2_model_chunks4.log
