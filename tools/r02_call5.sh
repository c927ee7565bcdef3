#!/bin/bash
timeout 1500 python -m pytest tests/test_gpu_knn.py tests/test_gpu_configs.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -8
PROBE_ITERS=3 timeout 300 python tools/knn_probe.py 2048 2048 128 91 303104 2>&1 | grep iter | sed -e 's/stats.*accepted\/row/acc\/row/' | cut -c1-260
PROBE_ITERS=2 timeout 300 python tools/knn_probe.py 2048 2048 128 91 524288 2>&1 | grep "iter 1" | sed -e 's/stats.*accepted\/row/acc\/row/' | cut -c1-260
SPHB_LIB=$PWD/sph_b200/lib/libsphb200_prof.so SPHB_KNN_PROF=1 SPHB_KNN_TIMING=1 PROBE_ITERS=1 timeout 300 python tools/knn_probe.py 2048 2048 128 91 303104 2>&1 | grep -E "knn prof|knn timing" | cut -c1-400
