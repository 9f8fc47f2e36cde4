#!/bin/bash
# splat iteration: parity subset, then softsplat / two-frame timings, tile gather against the plain gather
T=${1:-si}; mkdir -p gpurun_out
timeout -s KILL 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_edge_cases.py -q -m gpu -x -k "splat or fwarp or two_frame or mframes" > gpurun_out/${T}_tests.log 2>&1; echo "tests rc=$?"; tail -15 gpurun_out/${T}_tests.log
for g in tile plain; do
  for fl in smooth gentle iid; do
    VFI_SPLAT_GATHER=$g timeout -s KILL 100 python tools/op_bench.py --op splat --C 64 --level 0 --flow $fl 2>&1 | tail -1 | sed "s/^/$g /"
  done
  VFI_SPLAT_GATHER=$g timeout -s KILL 100 python tools/op_bench.py --op twoframe --C 64 --level 0 --flow smooth 2>&1 | tail -1 | sed "s/^/$g /"
  VFI_SPLAT_GATHER=$g timeout -s KILL 100 python tools/op_bench.py --op splat --C 16 --level 0 --flow smooth 2>&1 | tail -1 | sed "s/^/$g /"
done 2>&1 | tee gpurun_out/${T}_op.log
