#!/bin/bash
# tests, C5/C3 probes, ncu: DRAM/L2 bytes of one full C4 sweep launch + --set full of a 4-wave slab
python -m pytest tests -m gpu -q 2>&1 | tail -15 > gpurun_out/s14_pytest.log; tail -3 gpurun_out/s14_pytest.log
PROBE_ITERS=2 timeout 600 python tools/knn_probe.py 1024 1024 1000 91 151552 2>&1 | tail -2 > gpurun_out/s14_c5_probe.log; cat gpurun_out/s14_c5_probe.log | cut -c1-700
PROBE_ITERS=2 timeout 600 python tools/knn_probe.py 1024 1024 32 91 2>&1 | tail -1 > gpurun_out/s14_c3_probe.log; cat gpurun_out/s14_c3_probe.log | cut -c1-700
PROBE_ITERS=2 timeout 600 python tools/knn_probe.py 1024 1024 200 91 303104 2>&1 | tail -1 > gpurun_out/s14_d200_probe.log; cat gpurun_out/s14_d200_probe.log | cut -c1-700
PROBE_ITERS=1 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,lts__t_sectors_srcunit_tex_op_read.sum,sm__inst_executed_pipe_tensor.sum,sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_elapsed,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed,gpu__time_duration.sum,sm__cycles_elapsed.max,smsp__inst_executed.sum,l1tex__data_bank_conflicts_pipe_lsu.sum \
  --clock-control none -k regex:knn_sweep_resident -c 1 --csv --log-file gpurun_out/s14_sweep_c4_bytes.csv python tools/knn_probe.py 2048 2048 128 91 > gpurun_out/s14_ncu_bytes.log 2>&1; tail -2 gpurun_out/s14_ncu_bytes.log | cut -c1-400
PROBE_ITERS=1 ncu --set full --clock-control none --import-source on -k regex:knn_sweep_resident -c 1 -o gpurun_out/s14_sweep_lean_full -f python tools/knn_probe.py 2048 2048 128 91 151552 > gpurun_out/s14_ncu_full.log 2>&1; tail -2 gpurun_out/s14_ncu_full.log | cut -c1-400
ls -la gpurun_out/
