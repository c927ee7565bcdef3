#!/bin/bash
# One evidence pass on a B200: parity suite, bench line (+ reference arm), smoke, ncu launch list (after a plain run).
# Usage: gpurun --timeout 2000 -- bash tools/gpu_evidence.sh ; outputs land in gpurun_out/ev_*
timeout 1200 python -m pytest tests -m gpu -q 2>&1 | tail -4 > gpurun_out/ev_pytest.log; tail -2 gpurun_out/ev_pytest.log
python bench.py --steps 3 --warmup 3 > gpurun_out/ev_bench_c4.json 2> gpurun_out/ev_bench_c4.err; head -c 1200 gpurun_out/ev_bench_c4.json; echo
python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/ev_ref_c4.json 2> gpurun_out/ev_ref_c4.err; head -c 600 gpurun_out/ev_ref_c4.json; echo
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/ev_launches.csv python bench.py --steps 1 --warmup 0 --e2e-steps 0 --no-cpu-baseline --no-level1 > gpurun_out/ev_ncu_bench.log 2>&1; tail -1 gpurun_out/ev_ncu_bench.log | head -c 150; echo
