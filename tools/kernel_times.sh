#!/bin/bash
# Per-kernel durations + instruction counts of a command (ncu, cold cache): bash tools/kernel_times.sh <tag> <command...>
TAG=$1; shift
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none --csv --log-file gpurun_out/${TAG}_ktimes.csv "$@" > gpurun_out/${TAG}_ktimes.log 2>&1
echo rc=$?
