#!/bin/bash
# A/B of one env knob on the window bench, interleaved on one box:  KNOB=VDA_LN_REVERSE A=0 B=1 bash scripts/ab_env.sh
mkdir -p gpurun_out
OUT=gpurun_out/${TAG:-ab}_${KNOB}.log
: > $OUT
for rep in 1 2; do
  for v in $A $B; do
    env $KNOB=$v python bench.py --no-extras --no-cpu-baseline --steps 10 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline())
h=d.get('hbm_kernels',{})
print('$KNOB=$v', 'ms_per_step %.3f' % d['ms_per_step'], 'value %.1f' % d['value'], 'ln %.3f ms' % h.get('layernorm',{}).get('ms_per_step',0), 'gemm %.3f' % d['breakdown_ms_per_step']['gemm'], 'sm_mhz', d['clocks']['sm_mhz'])
" | tee -a $OUT
  done
done
