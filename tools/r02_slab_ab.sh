#!/bin/bash
# SLAB flavour of createSimilarities on a level too large for shared-memory accumulators (768 x 768 x 128 stack, level 1 = 147 456
# components, 1.66e11 products): reductions + touched bitmap (default) against returning atomics + first-touch list (round 1)
for bm in 1 0; do
  echo "== bitmap $bm"
  SPHB_LIB=$PWD/sph_b200/lib/libsphb200_prof.so SPHB_SIMS_BITMAP=$bm timeout 600 python tools/level1_c4.py --shape 768,768,128,32,1004 --parity-rows 64 2>&1 | tail -1 | python -c "
import json,sys
j=json.loads(sys.stdin.read().strip().splitlines()[-1])
print({k: j[k] for k in ('merged_nnz','d1_products','level1_prob_dist_ms')}, j['level1_stages_ms'], j.get('parity'))
"
done
