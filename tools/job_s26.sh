#!/bin/bash
run() { echo "== $1"; shift; env "$@" PROBE_ITERS=3 timeout 300 python tools/knn_probe.py 2048 2048 128 91 303104 2>&1 | tail -1 | grep -o "timing.*"; }
for rep in 1 2; do
run "pf 0" X=1
run "pf 8" SPHB_KNN_L2PF=8
run "pf 32" SPHB_KNN_L2PF=32
run "pf 128" SPHB_KNN_L2PF=128
run "pf 32 stagger 1" SPHB_KNN_L2PF=32 SPHB_KNN_STAGGER=1
run "pf 0 stagger 1" SPHB_KNN_STAGGER=1
done
SPHB_KNN_L2PF=32 timeout 300 python -m pytest tests/test_gpu_knn.py -x -q -k "tc_matches or hyperspectral" 2>&1 | tail -2
