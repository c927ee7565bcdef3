#!/bin/bash
run() { echo "== $1"; shift; env "$@" PROBE_ITERS=2 timeout 300 python tools/knn_probe.py 2048 2048 128 91 303104 2>&1 | grep "iter 1" | sed -e 's/stats.*accepted\/row/acc\/row/' | cut -c1-250; }
run "step cap (default)" X=1
run "step 256" SPHB_KNN_COMPACT_STEP=256
run "step 128" SPHB_KNN_COMPACT_STEP=128
run "step 64" SPHB_KNN_COMPACT_STEP=64
run "step 32" SPHB_KNN_COMPACT_STEP=32
run "step 64 stagger 2" SPHB_KNN_COMPACT_STEP=64 SPHB_KNN_STAGGER=2
run "pace 128" SPHB_KNN_PACE=128
run "pace 512" SPHB_KNN_PACE=512
