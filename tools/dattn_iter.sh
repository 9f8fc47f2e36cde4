#!/bin/bash
# kernel iteration loop for the fused attention: parity subset, op timings, optional ncu capture
#   bash tools/dattn_iter.sh <tag> [ncu]
T=${1:-it}; mkdir -p gpurun_out
timeout -s KILL 600 python -m pytest tests/test_gpu_parity_dat.py -q -m gpu -s -k "dattn or dat_block or tiled or full_width" > gpurun_out/${T}_tests.log 2>&1; echo "tests rc=$?"; grep -v "^$" gpurun_out/${T}_tests.log | tail -12
for lv in 1 2; do timeout -s KILL 200 python tools/op_bench.py --op dattn --level $lv --layout cl; done 2>&1 | tee gpurun_out/${T}_op.log
timeout -s KILL 200 python tools/op_bench.py --op dattn --level 1 --layout cl --flow iid 2>&1 | tee -a gpurun_out/${T}_op.log
timeout -s KILL 300 python bench.py --no-cpu --no-micro --no-e2e --steps 5 > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?"
python - <<PY
import json
d=json.load(open('gpurun_out/${T}_bench.json'))
print({k:round(v['ms_per_step'],3) for k,v in d['vfi_ops_ms'].items()}); print('dattn lv1 ms', d['roofline']['ms_per_launch'], 'step ms', d['ms_per_step'])
PY
if [ "${2:-}" = "ncu" ]; then bash tools/ncu_op.sh ${T}_ncu dattn_win --op dattn --level 1 --layout cl; fi
