# two-frame splat (a4): scatter (fp32 atomics) vs gather route, 1080p and 1/8 res
for fl in gentle smooth; do for c in 64 16; do for a in scatter gather; do
  VFI_SPLAT_ALGO=$a python tools/op_bench.py --op twoframe --level 0 --C $c --flow $fl --iters 10 2>&1 | tail -1 | cut -c1-260
done; done; done
