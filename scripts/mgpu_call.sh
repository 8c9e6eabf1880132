#!/bin/bash
# Multi-GPU validation under `gpurun --gpus N`: sharded-driver parity (NCCL), two devices in one process, and the
# headline bench with the sharded 1024-frame video inside, launched exactly as the driver launches it.
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
N=${N:-2}
TAG=${TAG:-r02}
nvidia-smi -L | tee gpurun_out/${TAG}_mgpu_gpus.txt
( time timeout -s KILL 900 python -m pytest tests/test_multigpu_gpu.py -x -q -m gpu -s ) > gpurun_out/${TAG}_mgpu_pytest.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/${TAG}_mgpu_pytest.log
( time timeout -s KILL 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
    bench.py --gpus $N --steps ${STEPS:-6} --warmup 3 ) > gpurun_out/${TAG}_bench_${N}gpu.json 2> gpurun_out/${TAG}_bench_${N}gpu.err; echo "bench rc=$?"
tail -3 gpurun_out/${TAG}_bench_${N}gpu.err
python - <<PY
import json
d=json.loads(open("gpurun_out/${TAG}_bench_${N}gpu.json").read().strip().split("\n")[-1])
print("value", d["value"], "e2e", d["e2e"]["value"], "p50", d["p50_window_latency_ms"])
print(json.dumps(d.get("extra"), indent=1)[:3000])
PY
