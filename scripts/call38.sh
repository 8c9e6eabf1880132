#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
O=gpurun_out
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 8000 -c 620 --csv --log-file $O/c38_stream_vits.csv \
  python scripts/stream_bench.py --arch vits --frames 50 > $O/c38_ncu.log 2>&1
python scripts/launch_summary.py $O/c38_stream_vits.csv x > $O/c38_stream_vits.txt; head -24 $O/c38_stream_vits.txt
