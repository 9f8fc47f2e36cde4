#!/bin/bash
# first GPU check of round 2: new parity tests, op timings of the two attention kernels, full suite, bench
set -u
mkdir -p gpurun_out
T=${1:-r3b}
timeout -s KILL 600 python -m pytest tests/test_gpu_parity_dat.py -q -m gpu -s -k "dattn or dat_block or tiled or full_width" > gpurun_out/${T}_newtests.log 2>&1; echo "new tests rc=$?"; tail -25 gpurun_out/${T}_newtests.log
for lv in 1 2; do
  timeout -s KILL 200 python tools/op_bench.py --op dattn --level $lv --layout cl >> gpurun_out/${T}_op.log 2>&1
  VFI_DATTN_ALGO=gather timeout -s KILL 200 python tools/op_bench.py --op dattn --level $lv --layout cl >> gpurun_out/${T}_op.log 2>&1
done
timeout -s KILL 200 python tools/op_bench.py --op dattn --level 1 --layout cl --flow iid >> gpurun_out/${T}_op.log 2>&1
cat gpurun_out/${T}_op.log
timeout -s KILL 900 python -m pytest tests -q -m gpu > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/${T}_pytest.log
timeout -s KILL 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/${T}_smoke.log
timeout -s KILL 600 python bench.py --no-cpu > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?"; cut -c1-1800 gpurun_out/${T}_bench.json; tail -3 gpurun_out/${T}_bench.err
