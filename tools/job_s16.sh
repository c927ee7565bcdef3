#!/bin/bash
timeout 300 python -m pytest tests/test_gpu_knn.py tests/test_gpu_graph.py -x -q 2>&1 | tail -4
run() { echo "== $1"; shift; env "$@" SPHB_LIB=$PWD/sph_b200/lib/libsphb200_prof.so SPHB_KNN_PROF=1 PROBE_ITERS=2 timeout 300 python tools/knn_probe.py 2048 2048 128 91 151552 2>&1 | grep -E "passes 1|iter 1" | tail -2 | cut -c1-420; }
run "pair" X=1
run "pair skip-epi" SPHB_KNN_DEBUG_SKIP_EPI=1
run "pair skip-epi local-tempty" SPHB_KNN_DEBUG_SKIP_EPI=1 SPHB_KNN_PAIR_DEBUG=1
run "pair skip-epi solo-tx" SPHB_KNN_DEBUG_SKIP_EPI=1 SPHB_KNN_PAIR_DEBUG=2
run "pair skip-epi no-B" SPHB_KNN_DEBUG_SKIP_EPI=1 SPHB_KNN_PAIR_DEBUG=4
run "pair skip-epi no-B local-tempty" SPHB_KNN_DEBUG_SKIP_EPI=1 SPHB_KNN_PAIR_DEBUG=5
run "single skip-epi" SPHB_KNN_DEBUG_SKIP_EPI=1 SPHB_KNN_NO_PAIR=1
