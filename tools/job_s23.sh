#!/bin/bash
timeout 1200 python -m pytest tests -m gpu -q 2>&1 | tail -6 > gpurun_out/s23_pytest.log; tail -3 gpurun_out/s23_pytest.log
python bench.py --steps 2 --warmup 3 > gpurun_out/s23_bench_c4.json 2> gpurun_out/s23_bench_c4.err; head -c 3000 gpurun_out/s23_bench_c4.json; echo
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/s23_launches.csv python bench.py --steps 1 --warmup 0 --e2e-steps 0 --no-cpu-baseline --no-level1 > gpurun_out/s23_ncu_bench.log 2>&1; tail -1 gpurun_out/s23_ncu_bench.log | head -c 200; echo
PROBE_ITERS=1 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,sm__inst_executed_pipe_tensor.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed,gpu__time_duration.sum,sm__cycles_elapsed.max,smsp__inst_executed.sum \
  --clock-control none -k regex:knn_sweep_resident --launch-skip 1 -c 1 --csv --log-file gpurun_out/s23_sweep_c4_main_bytes.csv python tools/knn_probe.py 2048 2048 128 91 > gpurun_out/s23_ncu_bytes.log 2>&1; tail -1 gpurun_out/s23_ncu_bytes.log | cut -c1-200
PROBE_ITERS=1 ncu --set full --clock-control none --import-source on -k regex:knn_sweep_resident --launch-skip 1 -c 1 -o gpurun_out/s23_sweep_full -f python tools/knn_probe.py 2048 2048 128 91 303104 > gpurun_out/s23_ncu_full.log 2>&1; tail -1 gpurun_out/s23_ncu_full.log | cut -c1-200
