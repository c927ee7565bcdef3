#!/bin/bash
run() { echo "== $1"; shift; env "$@" PROBE_ITERS=3 timeout 300 python tools/knn_probe.py 2048 2048 128 91 303104 2>&1 | tail -1 | grep -o "timing.*" | cut -c1-120; }
for rep in 1 2; do
run "default" X=1
run "spin" SPHB_LIB=$PWD/sph_b200/lib/libsphb200_spin.so
run "chunk-major" SPHB_KNN_CHUNK_MAJOR=1
run "stagger 2" SPHB_KNN_STAGGER=2
run "lead 4096" SPHB_KNN_LEAD_ROWS=4096
done
