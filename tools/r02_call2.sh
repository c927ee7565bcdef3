#!/bin/bash
# Round 2, second GPU call: new sims kernel + fused level pipeline tests, level-1 probe (after), sweep cycle counters.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_level.py tests/test_gpu_graph.py tests/test_gpu_yhierarchy.py -m gpu -x -q 2>&1 | tail -15
PROBE_REPS=2 timeout 300 python tools/level1_probe.py > gpurun_out/r02_level1_probe_after.json 2> gpurun_out/r02_level1_probe_after.err && cat gpurun_out/r02_level1_probe_after.json &&
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r02_level1_launches_after.csv python tools/level1_probe.py > gpurun_out/r02_level1_ncu_after.log 2>&1
tail -2 gpurun_out/r02_level1_ncu_after.log | head -c 300; echo
SPHB_LIB=$PWD/sph_b200/lib/libsphb200_prof.so SPHB_KNN_PROF=1 PROBE_ITERS=2 timeout 300 python tools/knn_probe.py 2048 2048 128 91 303104 2>&1 | grep -E "knn prof|iter" | tee gpurun_out/r02_sweep_prof.txt
