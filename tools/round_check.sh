#!/bin/bash
# Round check on the GPU box: GPU tests, smoke, both bench arms, launch list + one full ncu capture.
#   bash tools/round_check.sh <tag>
set -u
TAG=${1:-r1h}
mkdir -p gpurun_out
timeout -s KILL 900 python -m pytest tests -x -q -m gpu > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${TAG}_pytest.log
timeout -s KILL 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/${TAG}_smoke.log
timeout -s KILL 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/${TAG}_bench_ref.json 2> gpurun_out/${TAG}_bench_ref.err; echo "ref rc=$?"
timeout -s KILL 900 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"
cat gpurun_out/${TAG}_bench_ref.json; cut -c1-1500 gpurun_out/${TAG}_bench.json
timeout -s KILL 900 bash profiles/run_ncu.sh ${TAG} dattn_ 2 1
