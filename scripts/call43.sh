#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
O=gpurun_out
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 3000 -c 1200 --csv --log-file $O/c43_video_launches.csv \
  python scripts/video_bench.py --frames 256 --reps 1 --result-rank all > $O/c43_ncu.log 2>&1
echo rc=$?; tail -2 $O/c43_ncu.log | cut -c1-200
python scripts/launch_summary.py $O/c43_video_launches.csv > $O/c43_video_launches.txt; head -40 $O/c43_video_launches.txt
