#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
O=gpurun_out
timeout -s KILL 600 python -m pytest tests/ -x -q -m gpu -k "stream" 2>&1 | tail -3 | tee $O/c33_tests.log
for v in 1 0; do for a in vitl vitb vits; do VDA_TSTREAM=$v timeout 300 python scripts/stream_bench.py --arch $a --frames 120 | cut -c1-150 | sed "s/^/TSTREAM=$v /" | tee -a $O/c33_stream.log; done; done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 13500 -c 900 --csv --log-file $O/c33_stream_launches.csv \
  python scripts/stream_bench.py --arch vitl --frames 50 > $O/c33_ncu.log 2>&1
python scripts/launch_summary.py $O/c33_stream_launches.csv > $O/c33_stream_launches.txt; head -12 $O/c33_stream_launches.txt
