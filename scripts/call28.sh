#!/bin/bash
# per-kernel GPU durations of graph-replayed streaming frames (vitl): ncu launch list of the last ~2 frames
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
O=gpurun_out
timeout 300 python scripts/stream_bench.py --arch vitl --frames 50 | cut -c1-200
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 13500 -c 900 --csv --log-file $O/c28_stream_launches.csv \
  python scripts/stream_bench.py --arch vitl --frames 50 > $O/c28_ncu.log 2>&1
echo rc=$?; tail -2 $O/c28_ncu.log | cut -c1-200
python scripts/launch_summary.py $O/c28_stream_launches.csv per-launch > $O/c28_stream_launches.txt; head -30 $O/c28_stream_launches.txt
