#!/bin/bash
timeout 300 python -m pytest tests/test_gpu_knn.py -x -q -k "tc_matches_oracle" 2>&1 | tail -15 > gpurun_out/s15_pair_first.log; echo "rc=$?"; tail -5 gpurun_out/s15_pair_first.log
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -15 > gpurun_out/s15_pytest.log; tail -4 gpurun_out/s15_pytest.log
run() { echo "== $1"; shift; env "$@" PROBE_ITERS=3 timeout 300 python tools/knn_probe.py 2048 2048 128 91 151552 2>&1 | tail -1 | grep -o "timing.*"; }
run "pair" X=1
run "single" SPHB_KNN_NO_PAIR=1
run "pair" X=1
run "single" SPHB_KNN_NO_PAIR=1
run "pair skip-epi" SPHB_KNN_DEBUG_SKIP_EPI=1
run "single skip-epi" SPHB_KNN_DEBUG_SKIP_EPI=1 SPHB_KNN_NO_PAIR=1
PROBE_ITERS=2 timeout 600 python tools/knn_probe.py 1024 1024 1000 91 151552 2>&1 | tail -1 | cut -c1-600
