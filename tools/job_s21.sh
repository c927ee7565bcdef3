#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_knn.py -x -q 2>&1 | tail -4
run() { echo "== $1"; shift; env "$@" PROBE_ITERS=3 timeout 300 python tools/knn_probe.py 2048 2048 128 91 151552 2>&1 | tail -1 | grep -o "resweep_rows.*"; }
run "ladder" X=1
run "ladder" X=1
prof() { echo "== $1"; shift; env "$@" SPHB_LIB=$PWD/sph_b200/lib/libsphb200_prof.so SPHB_KNN_PROF=1 PROBE_ITERS=2 timeout 300 python tools/knn_probe.py 2048 2048 128 91 151552 2>&1 | grep -E "passes 1" | tail -1 | cut -c1-420; }
prof "ladder" X=1
PROBE_ITERS=1 ncu --set full --clock-control none --import-source on -k regex:knn_sweep_resident --launch-skip 1 -c 1 -o gpurun_out/s21_sweep_full -f python tools/knn_probe.py 2048 2048 128 91 151552 > gpurun_out/s21_ncu_full.log 2>&1; tail -1 gpurun_out/s21_ncu_full.log | cut -c1-300
PROBE_ITERS=1 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,sm__inst_executed_pipe_tensor.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed,gpu__time_duration.sum,sm__cycles_elapsed.max,smsp__inst_executed.sum \
  --clock-control none -k regex:knn_sweep_resident --launch-skip 1 -c 1 --csv --log-file gpurun_out/s21_sweep_c4_main_bytes.csv python tools/knn_probe.py 2048 2048 128 91 > gpurun_out/s21_ncu_bytes.log 2>&1; tail -1 gpurun_out/s21_ncu_bytes.log | cut -c1-300
