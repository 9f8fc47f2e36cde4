#!/bin/bash
# full round check: GPU tests, smoke, both bench arms
T=${1:-full}; mkdir -p gpurun_out
timeout -s KILL 1200 python -m pytest tests -q -m gpu -s > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"; grep -E "passed|failed|measured mode|impulse|tiled vs|re-basing" gpurun_out/${T}_pytest.log | tail -12
timeout -s KILL 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/${T}_smoke.log
timeout -s KILL 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${T}_bench_ref.json 2> gpurun_out/${T}_bench_ref.err; echo "ref rc=$?"; cut -c1-600 gpurun_out/${T}_bench_ref.json
SECONDS=0
timeout -s KILL 1500 python bench.py > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$? in ${SECONDS}s"; tail -5 gpurun_out/${T}_bench.err
python - <<PY
import json
d=json.load(open('gpurun_out/${T}_bench.json'))
for k in ('value','ms_per_step','e2e','roofline','roofline_conv3x3','dcndat_1080p','tiled_4k','configs0_448x256','e2e_parity','value_fp32_convs','torch_cuda_baseline','cpu_baseline'):
    print(k, json.dumps(d.get(k))[:700])
m=d.get('microbench',{})
for k in m:
    if k!='cases': print(k, json.dumps(m[k])[:1500])
PY
