#!/bin/bash
# One --set full capture of an op's kernels through tools/op_bench.py (run on the GPU box through gpurun).
#   bash tools/ncu_op_capture.sh <tag> <op> [kernel-regex] [extra op_bench args]
TAG=$1; OP=$2; KRE=${3:-.}; shift 3
CMD="python tools/op_bench.py --op $OP --iters 2 $*"
$CMD > gpurun_out/${TAG}_${OP}_plain.log 2>&1 || { echo plain failed; tail -5 gpurun_out/${TAG}_${OP}_plain.log; exit 1; }
cat gpurun_out/${TAG}_${OP}_plain.log
ncu --set full --clock-control none --import-source on -k regex:${KRE} -s ${NCU_SKIP:-3} -c ${NCU_COUNT:-1} -o gpurun_out/${TAG}_${OP} -f $CMD > gpurun_out/${TAG}_${OP}_ncu.log 2>&1
echo rc=$?
