#!/bin/bash
timeout 600 python -m pytest tests/test_gpu_knn.py -x -q 2>&1 | tail -3
run() { echo "== $1"; shift; env "$@" PROBE_ITERS=3 timeout 300 python tools/knn_probe.py 2048 2048 128 91 303104 2>&1 | tail -1 | grep -o "accepted/row.*"; }
for rep in 1 2; do
run "step=cap (default)" X=1
run "step 192" SPHB_KNN_COMPACT_STEP=192
run "step 128" SPHB_KNN_COMPACT_STEP=128
run "step 96" SPHB_KNN_COMPACT_STEP=96
run "step 64" SPHB_KNN_COMPACT_STEP=64
done
SPHB_KNN_COMPACT_STEP=128 timeout 600 python -m pytest tests/test_gpu_knn.py -x -q 2>&1 | tail -3
