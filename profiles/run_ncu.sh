#!/bin/bash
# Profiling recipe of this repo (run on the GPU box through gpurun; see /opt/skills/guides/B200_PROFILING.md).
#   bash profiles/run_ncu.sh <tag> [kernel-regex] [skip] [count]
# 1. plain run must exit 0; 2. launch list (gpu__time_duration, cold cache, serialised: compare SHARES);
# 3. one --set full capture of the kernel named by the regex.
set -u
TAG=${1:-r1}; KRE=${2:-dattn_kernel}; SKIP=${3:-0}; CNT=${4:-3}
CMD="python bench.py --steps 1 --warmup 3 --no-micro --no-cpu --no-e2e --graph 0 ${BENCH_EXTRA:-}"
mkdir -p gpurun_out
$CMD > gpurun_out/${TAG}_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${TAG}_plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu_list.log 2>&1
echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:${KRE} -s ${SKIP} -c ${CNT} -o gpurun_out/${TAG}_${KRE} -f $CMD > gpurun_out/${TAG}_ncu_full.log 2>&1
echo "full capture rc=$?"
ls -la gpurun_out/
