#!/bin/bash
# A/B of the thread-group shape of the DENSE createSimilarities kernel (profile build reads SPHB_SIMS_CFG):
# 0 = class-partitioned barrier-free kernel (default), 10 = column-at-a-time kernel with 16 warps (1..4: its 8- and 4-warp shapes)
for cfg in 0 10; do
  echo "== cfg $cfg"
  SPHB_LIB=$PWD/sph_b200/lib/libsphb200_prof.so SPHB_SIMS_CFG=$cfg PROBE_REPS=2 timeout 300 python tools/level1_probe.py 2>&1 | python -c "
import json,sys
j=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('similarities_ms', j['fused_stage_ms_rep1']['similarities_ms'], 'fused==staged', j['fused_equals_staged_rep1'], 'p_nnz', j['p_nnz'], 'col len mean/max', j['sims_col_len_mean'], j['sims_col_len_max'])
"
done
