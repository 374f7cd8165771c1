#!/bin/bash
# ncu --set full capture of K2 on config 3 (1M segments vs the 256-OBB world).  Usage: bash profiles/prof_segments.sh [tag]
TAG=${1:-r1}
OUT=gpurun_out; mkdir -p $OUT
SMALL="python bench.py --workload segments --steps 32 --warmup 3 --no-graph --no-extras --no-cpu-baseline"
timeout 300 $SMALL > $OUT/${TAG}_seg_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_segments -s 5 -c 2 -f -o $OUT/${TAG}_k_segments $SMALL > $OUT/${TAG}_ncu_seg.log 2>&1
echo "ncu exit $?"
