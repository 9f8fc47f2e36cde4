#!/bin/bash
# Launch list with caches NOT flushed between kernels (closer to in-step durations than the default cold list).
#   bash tools/warm_launch_list.sh <tag> [extra bench args]
TAG=$1; shift
CMD="python bench.py --steps 1 --warmup 3 --no-micro --no-cpu --no-e2e --graph 0 $*"
ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none --csv --log-file gpurun_out/${TAG}_launches_warm.csv $CMD > gpurun_out/${TAG}_warm.log 2>&1
echo "rc=$?"
