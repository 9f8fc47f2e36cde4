#!/bin/bash
# kernel iteration loop for bwarp / splat: parity subset, op timings
T=${1:-wi}; mkdir -p gpurun_out
timeout -s KILL 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_edge_cases.py -q -m gpu -x > gpurun_out/${T}_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/${T}_tests.log
for fl in smooth gentle iid; do
  timeout -s KILL 100 python tools/op_bench.py --op bwarp --C 64 --level 0 --flow $fl
  timeout -s KILL 100 python tools/op_bench.py --op splat --C 64 --level 0 --flow $fl
done 2>&1 | tee gpurun_out/${T}_op.log
timeout -s KILL 100 python tools/op_bench.py --op bwarp --C 16 --level 0 --flow smooth | tee -a gpurun_out/${T}_op.log
