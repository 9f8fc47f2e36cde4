#!/bin/bash
# N-GPU check on one box: frame-pair sharded 1080p bench (weak scaling) and the 4K spatially tiled pair (strong scaling).
#   bash tools/multi_gpu_check.sh <tag> <N>
TAG=$1; N=$2
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
$TR bench.py --gpus $N --steps 10 --warmup 3 --no-micro --no-cpu > gpurun_out/${TAG}_bench_${N}gpu.json 2> gpurun_out/${TAG}_bench_${N}gpu.err; echo "bench rc=$?"
$TR bench.py --gpus $N --steps 10 --warmup 3 --height 2160 --width 3840 --tile-grid 2x4 --halo 64 > gpurun_out/${TAG}_bench_4k_tiled_${N}gpu.json 2> gpurun_out/${TAG}_tiled_${N}gpu.err; echo "tiled rc=$?"
cut -c1-400 gpurun_out/${TAG}_bench_${N}gpu.json; cut -c1-300 gpurun_out/${TAG}_bench_4k_tiled_${N}gpu.json
