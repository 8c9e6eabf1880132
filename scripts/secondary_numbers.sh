#!/bin/bash
# Refresh of the secondary numbers (BASELINE configs 3-5, FP32-accuracy mode, other archs) on one GPU.
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
O=gpurun_out
timeout -s KILL 600 python scripts/video_bench.py --frames 1024 --reps 3 2>$O/rf_video_vitl.err | grep '"workload"' | tail -n 1 > $O/rf_video_vitl.json; cut -c1-330 $O/rf_video_vitl.json
timeout -s KILL 600 python scripts/video_bench.py --arch vitb --frames 512 --height 720 --width 1280 --metric --reps 3 2>$O/rf_video_vitb.err | grep '"workload"' | tail -n 1 > $O/rf_video_vitb.json; cut -c1-330 $O/rf_video_vitb.json
for arch in vits vitl; do timeout -s KILL 300 python scripts/stream_bench.py --arch $arch 2>/dev/null | tail -n 1 > $O/rf_stream_$arch.json; cut -c1-200 $O/rf_stream_$arch.json; done
timeout -s KILL 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --precision fp32 2>/dev/null > $O/rf_bench_fp32.json; cut -c1-200 $O/rf_bench_fp32.json
for arch in vitb vits; do timeout -s KILL 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --arch $arch 2>/dev/null > $O/rf_bench_$arch.json; cut -c1-160 $O/rf_bench_$arch.json; done
