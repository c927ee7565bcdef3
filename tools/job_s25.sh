#!/bin/bash
timeout 300 python -m pytest tests/test_gpu_knn.py -x -q -k "async" 2>&1 | tail -4
timeout 600 python -m pytest tests/test_gpu_knn.py -x -q 2>&1 | tail -3
run() { echo "== $1"; shift; env "$@" PROBE_ITERS=3 timeout 300 python tools/knn_probe.py 2048 2048 128 91 303104 2>&1 | tail -1 | grep -o "accepted/row.*"; }
for rep in 1 2; do
run "sync (default)" X=1
run "async" SPHB_KNN_ASYNC=1
done
prof() { echo "== $1"; shift; env "$@" SPHB_LIB=$PWD/sph_b200/lib/libsphb200_prof.so SPHB_KNN_PROF=1 PROBE_ITERS=2 timeout 300 python tools/knn_probe.py 2048 2048 128 91 303104 2>&1 | grep -E "passes 1" | tail -1 | cut -c1-420; }
prof "sync" X=1
prof "async" SPHB_KNN_ASYNC=1
