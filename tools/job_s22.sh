#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_knn.py -x -q 2>&1 | tail -4
run() { echo "== $1"; shift; env "$@" PROBE_ITERS=3 timeout 300 python tools/knn_probe.py 2048 2048 128 91 303104 2>&1 | tail -1 | grep -o "accepted/row.*"; }
for rep in 1 2; do
run "pace 256 ladder" X=1
run "pace off ladder" SPHB_KNN_PACE=0
run "pace 256 noladder" SPHB_LIB=$PWD/sph_b200/lib/libsphb200_noladder.so
run "pace 64" SPHB_KNN_PACE=64
run "pace 1024" SPHB_KNN_PACE=1024
done
PROBE_ITERS=1 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed,gpu__time_duration.sum \
  --clock-control none -k regex:knn_sweep_resident --launch-skip 1 -c 1 --csv --log-file gpurun_out/s22_sweep_main_bytes.csv python tools/knn_probe.py 2048 2048 128 91 303104 > gpurun_out/s22_ncu_bytes.log 2>&1; grep -E "dram__bytes|duration|tensor" gpurun_out/s22_sweep_main_bytes.csv | cut -d, -f13-
