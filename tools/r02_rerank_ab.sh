#!/bin/bash
# A/B of the staged re-rank and the 1-product first stage of the overflow re-sweep (profile build: knobs from the environment).
run() { echo "== $1"; shift; env "$@" SPHB_LIB=$PWD/sph_b200/lib/libsphb200_prof.so PROBE_ITERS=2 SPHB_KNN_TIMING=1 timeout 400 python tools/knn_probe.py 2048 2048 128 91 ${ROWS:-303104} 2>&1 | grep -E "iter 1|knn timing" | sed -e 's/stats.*accepted\/row/acc\/row/' | cut -c1-300; }
run "staged rerank (default)" X=1
run "old rerank" SPHB_KNN_RERANK_STAGED=0
run "staged + 1-product re-sweep stage" SPHB_KNN_RESWEEP_1PASS=1
ROWS=4194304
run "full C4: default" X=1
run "full C4: 1-product re-sweep stage" SPHB_KNN_RESWEEP_1PASS=1
run "full C4: old rerank" SPHB_KNN_RERANK_STAGED=0
