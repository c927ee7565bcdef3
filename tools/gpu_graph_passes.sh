#!/bin/bash
# First GPU call for the graph passes of csrc/graph_merge.cu (written after round 1's GPU budget was spent):
# hardware parity run of the emulation-verified tests, a timing probe at C2 size, an ncu launch list of the probe.
# Usage: gpurun --timeout 1500 -- bash tools/gpu_graph_passes.sh ; outputs land in gpurun_out/gp_*
timeout 900 python -m pytest tests/test_gpu_yhierarchy.py tests/test_gpu_zgraph.py -m gpu --runxfail -q 2>&1 | tail -15 > gpurun_out/gp_pytest.log; tail -3 gpurun_out/gp_pytest.log
timeout 600 python tools/graph_probe.py > gpurun_out/gp_probe.json 2> gpurun_out/gp_probe.err; head -c 2000 gpurun_out/gp_probe.json; echo
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 600 --csv --log-file gpurun_out/gp_launches.csv python tools/graph_probe.py 20000 31 > gpurun_out/gp_ncu.log 2>&1; tail -1 gpurun_out/gp_ncu.log | head -c 200; echo
