#!/bin/bash
timeout 600 python -m pytest tests/test_gpu_knn.py -x -q 2>&1 | tail -4
run() { echo "== $1"; shift; env "$@" PROBE_ITERS=3 timeout 300 python tools/knn_probe.py 2048 2048 128 91 151552 2>&1 | tail -1 | grep -o "compactions/row.*"; }
for rep in 1 2; do
run "pair select" X=1
run "single select" SPHB_KNN_NO_PAIR=1
run "pair fullsort" SPHB_LIB=$PWD/sph_b200/lib/libsphb200_fullsort.so
run "single fullsort" SPHB_KNN_NO_PAIR=1 SPHB_LIB=$PWD/sph_b200/lib/libsphb200_fullsort.so
done
prof() { echo "== $1"; shift; env "$@" SPHB_LIB=$PWD/sph_b200/lib/libsphb200_prof.so SPHB_KNN_PROF=1 PROBE_ITERS=2 timeout 300 python tools/knn_probe.py 2048 2048 128 91 151552 2>&1 | grep -E "passes 1" | tail -1 | cut -c1-420; }
prof "pair" X=1
prof "single" SPHB_KNN_NO_PAIR=1
