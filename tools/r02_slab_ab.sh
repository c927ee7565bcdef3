#!/bin/bash
# SLAB flavour of createSimilarities (768 x 768 x 128 stack, level 1 = 147 456 components, 1.66e11 products): number of CTAs
# (= live accumulator slabs of 590 KB each; 126 MB of L2 hold ~200) with the reductions + bitmap kernel (profile build).
# Measured: 296 CTAs (default, 2 per SM) 2.93 s, 148: 3.31 s, 74: 4.85 s, 444: 3.48 s, 592: 3.26 s -> the L2 reduction rate
# (~5.6e10 float reductions/s), not slab residency or occupancy, bounds the kernel.
for ctas in 296 148 74 444 592; do
  echo "== $ctas CTAs"
  SPHB_LIB=$PWD/sph_b200/lib/libsphb200_prof.so SPHB_SIMS_SLAB_CTAS=$ctas timeout 600 python tools/level1_c4.py --shape 768,768,128,32,1004 --parity-rows 0 2>&1 | tail -1 | python -c "
import json,sys
j=json.loads(sys.stdin.read().strip().splitlines()[-1])
print(j['level1_stages_ms']['similarities_ms'])
"
done
