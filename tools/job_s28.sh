#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_knn.py -x -q 2>&1 | tail -3
PROBE_ITERS=5 timeout 300 python tools/knn_probe.py 2048 2048 128 91 303104 2>&1 | grep -o "iter.*ms  alg\|resweep_rows.*" | cut -c1-260
PROBE_ITERS=3 timeout 300 python tools/knn_probe.py 2048 2048 128 91 524288 2>&1 | grep -o "iter.*ms  alg\|resweep_rows.*" | cut -c1-260
