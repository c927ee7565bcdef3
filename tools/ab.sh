#!/bin/bash
run() { echo "== $1"; shift; env "$@" PROBE_ITERS=3 python tools/knn_probe.py 2048 2048 128 91 151552 2>&1 | tail -1 | grep -o "sweep_ms[^,]*"; }
python -m pytest tests/test_gpu_knn.py -x -q 2>&1 | tail -3
for rep in 1 2; do
run "dual half-major" X=1
run "single chunk-major" SPHB_KNN_CHUNK_MAJOR=1
done
run "skip-epi dual" SPHB_KNN_DEBUG_SKIP_EPI=1
SPHB_KNN_TRACE_T0=20000 SPHB_KNN_TRACE_FILE=gpurun_out/trace.txt SPHB_LIB=$PWD/sph_b200/lib/libsphb200_prof.so SPHB_KNN_PROF=1 PROBE_ITERS=1 python tools/knn_probe.py 2048 2048 128 91 37888 2>&1 | grep "passes 1" | tail -1
