#!/bin/bash
# last GPU call of round 2: the new geodesic tests first (so a short budget still answers for them), then the rest of the -m gpu suite
mkdir -p gpurun_out
timeout 75 python -m pytest tests/test_zx_gpu_geodesic.py -v -m gpu --timeout 40 --timeout-method=thread -p no:cacheprovider 2>&1 | tee gpurun_out/r02i_gpu_geodesic.log | tail -12
timeout 115 python -m pytest tests -q -m gpu --ignore tests/test_zx_gpu_geodesic.py --timeout 60 --timeout-method=thread -p no:cacheprovider 2>&1 | tee gpurun_out/r02i_gpu_suite.log | tail -4
