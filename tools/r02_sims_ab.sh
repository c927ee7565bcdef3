#!/bin/bash
# createSimilarities at C2 level 1 (tools/level1_probe.py).  History of this A/B (one B200, ms for the similarities stage):
#   column-at-a-time kernel, block barrier per column (16 warps x 2 entries x 8 columns)        90.5
#     the same with 8 warps x 4 x 4 / 8 x 3 x 6 / 4 x 6 x 3 / 8 x 2 x 8                          108.6 / 114.7 / 140.5 / 140.6
#   class-partitioned kernel (rows j mod 16 -> 16 warps, no block barrier), per-warp sorted lists 85.2
#   + bottom-k candidates from a two-level histogram                                             56.0
#   + 8-byte (row, value) records                                                                52.3   <- the kernel in the tree
#   + one cp.async.bulk + mbarrier per segment into a per-warp byte ring                          139.3  (deleted)
#   + 16-byte cp.async copies into a three-stage per-warp buffer                                  80.6   (deleted)
SPHB_LIB=$PWD/sph_b200/lib/libsphb200_prof.so PROBE_REPS=2 timeout 300 python tools/level1_probe.py 2>&1 | python -c "
import json,sys
j=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('similarities_ms', j['fused_stage_ms_rep1']['similarities_ms'], 'fused==staged', j['fused_equals_staged_rep1'], 'p_nnz', j['p_nnz'])
"
