#!/bin/bash
# Per-kernel GPU durations of graph-replayed streaming frames: ncu launch list of the last frames of a short stream.
# (This is how the streaming cache attention was found to take 2.3 of a vitl frame's 6.0 ms in round 2.)
#   ARCH=vitl SKIP=13500 N=900 TAG=r02 bash scripts/stream_launch_list.sh      (vits: SKIP=8000 N=620)
mkdir -p gpurun_out
ARCH=${ARCH:-vitl}; SKIP=${SKIP:-13500}; N=${N:-900}; TAG=${TAG:-r02}
O=gpurun_out
timeout 300 python scripts/stream_bench.py --arch $ARCH --frames 50 | cut -c1-200 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip $SKIP -c $N --csv \
  --log-file $O/${TAG}_stream_${ARCH}_launches.csv python scripts/stream_bench.py --arch $ARCH --frames 50 > $O/${TAG}_stream_ncu.log 2>&1
python scripts/launch_summary.py $O/${TAG}_stream_${ARCH}_launches.csv per-launch > $O/${TAG}_stream_${ARCH}_launches.txt
head -24 $O/${TAG}_stream_${ARCH}_launches.txt
