#!/bin/bash
# A/B of conv3x3 variants: each argument is either a variant name (variants/libvfi_<name>.so) or ENV=VALUE settings for the
# in-tree library (e.g. "VFI_CONV_OSTG=0"); "default" = in-tree library as is.  Parity tests, per-shape table head, bench step.
set -u
mkdir -p gpurun_out
for n in "$@"; do
  unset VFI_B200_LIB VFI_CONV_OSTG VFI_CONV_RSTG VFI_CONV_STAGES VFI_DAT_BATCH_FRAMES VFI_CONV_DBG
  if [ -f "$PWD/variants/libvfi_$n.so" ]; then export VFI_B200_LIB=$PWD/variants/libvfi_$n.so; elif [ "$n" != default ]; then export $n; fi
  echo "=== $n"
  timeout -s KILL 300 python -m pytest tests/test_gpu_conv.py -q -m gpu -x 2>&1 | tail -3
  timeout -s KILL 200 python tools/conv_table.py 2>&1 | head -${CONV_AB_ROWS:-8}
  timeout -s KILL 300 python bench.py --no-cpu --no-legs --no-micro --steps 10 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.readline()); print('ms_per_step', d['ms_per_step'], 'value', d['value'], 'e2e', d['e2e']['value'])"
done 2>&1 | tee gpurun_out/conv_ab.log
