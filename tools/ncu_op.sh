#!/bin/bash
# one --set full capture of a kernel of tools/op_bench.py:  bash tools/ncu_op.sh <tag> <kernel regex> <op_bench args...>
TAG=$1; KRE=$2; shift 2
mkdir -p gpurun_out
CMD="python tools/op_bench.py --iters 2 $*"
$CMD > gpurun_out/${TAG}_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${TAG}_plain.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:${KRE} -s 3 -c 1 -o gpurun_out/${TAG} -f $CMD > gpurun_out/${TAG}_ncu.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/${TAG}_ncu.log
