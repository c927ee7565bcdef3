#!/bin/bash
# C4 bench line at N=1 + launch list + DRAM bytes of the main sweep launch
mkdir -p gpurun_out
python bench.py --steps 3 --warmup 3 > gpurun_out/r02f_bench_c4_n1.json 2> gpurun_out/r02f_bench_c4_n1.err; tail -3 gpurun_out/r02f_bench_c4_n1.err; python - <<'PY'
import json
j=json.load(open('gpurun_out/r02f_bench_c4_n1.json'))
print({k:j[k] for k in ('value','ms_per_step','clocks','gpu_launches','parity','cpu_baseline')})
print('e2e',j['e2e']); print('roofline',j['roofline'])
for p in j['graph_pass_roofline']: print(p['pass'],round(p['kernel_ms'],3),'ms',round(p['achieved'],1),'GB/s',round(p['frac'],3))
print('level1',{k:v for k,v in j['level1_p_matrix'].items() if k!='workload'})
PY
python bench.py --steps 1 --warmup 1 --e2e-steps 1 --no-cpu-baseline --no-level1 --parity-rows 0 > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed --clock-control none -c 200 --csv --log-file gpurun_out/r02f_ncu_launches_c4_step.csv python bench.py --steps 1 --warmup 1 --e2e-steps 1 --no-cpu-baseline --no-level1 --parity-rows 0 > gpurun_out/ncu_bench.log 2>&1; tail -1 gpurun_out/ncu_bench.log | head -c 200; echo
PROBE_REPS=1 timeout 300 python tools/level1_probe.py > gpurun_out/plain2.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r02f_level1_launches.csv python tools/level1_probe.py > gpurun_out/r02f_level1_ncu.log 2>&1; tail -1 gpurun_out/r02f_level1_ncu.log | head -c 300; echo
