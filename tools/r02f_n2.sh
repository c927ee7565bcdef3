#!/bin/bash
# 2-GPU evidence: the single-process multi-device entry point on two real devices (persistent worker threads)
timeout 300 python -m pytest tests/test_gpu_multi.py -q -m gpu 2>&1 | tail -2
timeout 600 python tools/multi_probe.py C4 2>&1 | tail -1 | tee gpurun_out/r02g_multi_probe_c4_g2.json
