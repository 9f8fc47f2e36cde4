#!/bin/bash
# DCN window kernel iteration: parity, then BASELINE configs[2] timings per algo with and without frame-leaving flows
set -u
mkdir -p gpurun_out
T=${1:-dcn}
timeout -s KILL 300 python -m pytest tests/test_gpu_parity_dat.py tests/test_gpu_parity.py -q -m gpu -k "dcn" > gpurun_out/${T}_tests.log 2>&1; echo "tests rc=$?"; tail -5 gpurun_out/${T}_tests.log
for algo in window simt auto; do
  for o in 0.02 0.0 0.10; do
    VFI_DCN_ALGO=$algo timeout -s KILL 200 python tools/op_bench.py --op dcn --outliers $o 2>&1 | tail -1
  done
done | tee gpurun_out/${T}_op.log
