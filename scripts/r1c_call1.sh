#!/bin/bash
# GPU call 1 of this session: parity of the new GEMM tail split-K + PDL launches, then A/B timings.
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
O=gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > $O/c1_smi.log
timeout -s KILL 400 python -m pytest tests/test_gemm_gpu.py -q -m gpu -x > $O/c1_gemm.log 2>&1; echo "gemm rc=$?" | tee -a $O/c1_rc.log
timeout -s KILL 900 python -m pytest tests/test_model_gpu.py -q -m gpu -x -s > $O/c1_model.log 2>&1; echo "model rc=$?" | tee -a $O/c1_rc.log
tail -n 4 $O/c1_gemm.log $O/c1_model.log
timeout -s KILL 300 python scripts/bench_ops.py proj fc2 --env VDA_GEMM_KSPLIT=0,1 > $O/c1_ksplit.log 2>&1; echo "ksplit rc=$?" | tee -a $O/c1_rc.log
cat $O/c1_ksplit.log
for pdl in 0 1; do
  VDA_PDL=$pdl timeout -s KILL 400 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > $O/c1_bench_pdl$pdl.json 2> $O/c1_bench_pdl$pdl.err; echo "bench pdl=$pdl rc=$?" | tee -a $O/c1_rc.log
  VDA_PDL=$pdl timeout -s KILL 300 python scripts/stream_bench.py --arch vits > $O/c1_stream_vits_pdl$pdl.json 2> $O/c1_stream_pdl$pdl.err; echo "stream pdl=$pdl rc=$?" | tee -a $O/c1_rc.log
done
VDA_PDL=1 VDA_GEMM_KSPLIT=0 timeout -s KILL 400 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > $O/c1_bench_pdl1_ks0.json 2>> $O/c1_bench_pdl1.err
cut -c1-260 $O/c1_bench_pdl0.json $O/c1_bench_pdl1.json $O/c1_bench_pdl1_ks0.json
cut -c1-200 $O/c1_stream_vits_pdl0.json $O/c1_stream_vits_pdl1.json
timeout -s KILL 300 python scripts/prof_launches.py vitl > $O/c1_prof_launches.txt 2>&1; echo "prof rc=$?" | tee -a $O/c1_rc.log
head -45 $O/c1_prof_launches.txt
