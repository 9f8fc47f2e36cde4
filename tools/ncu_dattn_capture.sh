TAG=$1
CMD="python bench.py --steps 1 --warmup 3 --no-micro --no-cpu --no-e2e --graph 0"
ncu --set full --clock-control none --import-source on -k regex:dattn_ -s 2 -c 1 -o gpurun_out/${TAG}_dattn -f $CMD > gpurun_out/${TAG}_ncu_full.log 2>&1
echo rc=$?
