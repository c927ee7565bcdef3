#!/bin/bash
timeout 1200 python -m pytest tests -m gpu -q 2>&1 | tail -4 > gpurun_out/s29_pytest.log; tail -2 gpurun_out/s29_pytest.log
python bench.py --steps 3 --warmup 3 > gpurun_out/s29_bench_c4.json 2> gpurun_out/s29_bench_c4.err; head -c 1200 gpurun_out/s29_bench_c4.json; echo
python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/s29_ref_c4.json 2> gpurun_out/s29_ref_c4.err; head -c 600 gpurun_out/s29_ref_c4.json; echo
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/s29_launches.csv python bench.py --steps 1 --warmup 0 --e2e-steps 0 --no-cpu-baseline --no-level1 > gpurun_out/s29_ncu_bench.log 2>&1; tail -1 gpurun_out/s29_ncu_bench.log | head -c 150; echo
