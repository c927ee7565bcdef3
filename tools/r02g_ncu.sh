#!/bin/bash
python bench.py --steps 1 --warmup 1 --e2e-steps 1 --no-cpu-baseline --no-level1 --parity-rows 0 > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed --clock-control none -c 200 --csv --log-file gpurun_out/r02g_ncu_launches_c4_step.csv python bench.py --steps 1 --warmup 1 --e2e-steps 1 --no-cpu-baseline --no-level1 --parity-rows 0 > gpurun_out/ncu_bench.log 2>&1; tail -1 gpurun_out/ncu_bench.log | head -c 300; echo
