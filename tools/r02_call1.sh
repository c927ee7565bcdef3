#!/bin/bash
# Round 2, first GPU call: UMMA operand/tile-shape probe, graph_merge tests without xfail, per-kernel ncu list of the level-1
# passes at C2 (the "before" numbers), fp64-pipe counters of the P-row kernel.
mkdir -p gpurun_out
timeout 300 tools/build/umma_probe 4096 > gpurun_out/r02_umma_probe.txt 2>&1; tail -45 gpurun_out/r02_umma_probe.txt
timeout 600 python -m pytest tests/test_gpu_zgraph.py tests/test_gpu_prow.py -m gpu -q 2>&1 | tail -3
timeout 300 python tools/level1_probe.py > gpurun_out/r02_level1_probe.json 2> gpurun_out/r02_level1_probe.err && cat gpurun_out/r02_level1_probe.json &&
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r02_level1_launches_before.csv python tools/level1_probe.py > gpurun_out/r02_level1_ncu.log 2>&1
tail -2 gpurun_out/r02_level1_ncu.log | head -c 300; echo
timeout 120 python tools/prow_probe.py 1048576 91 > gpurun_out/r02_prow_probe.txt 2>&1 && cat gpurun_out/r02_prow_probe.txt &&
timeout 600 ncu --set full --clock-control none -k regex:prow_graph_kernel -s 1 -c 1 -o gpurun_out/r02_prow_before python tools/prow_probe.py 1048576 91 > gpurun_out/r02_prow_ncu.log 2>&1
tail -2 gpurun_out/r02_prow_ncu.log | head -c 300; echo
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv
