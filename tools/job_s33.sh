#!/bin/bash
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -3
PROBE_ITERS=10 timeout 300 python tools/knn_probe.py 2048 2048 128 91 524288 2>&1 | grep -o "iter.*ms  alg\|'sweep_ms.*" | cut -c1-200
