#!/bin/bash
# same-box A/B of the DAT 1080p step: arguments as tools/conv_ab.sh; two rounds, 20 steps each
set -u
for round in 1 2; do
for n in "$@"; do
  unset VFI_B200_LIB VFI_CONV_OSTG VFI_CONV_RSTG VFI_CONV_STAGES VFI_DAT_BATCH_FRAMES VFI_CONV_DBG
  if [ -f "$PWD/variants/libvfi_$n.so" ]; then export VFI_B200_LIB=$PWD/variants/libvfi_$n.so; elif [ "$n" != default ]; then export $n; fi
  timeout -s KILL 300 python bench.py --no-cpu --no-legs --no-micro --steps 20 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.readline()); print('$n', 'ms_per_step', round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value'],2))"
done
done
