#!/bin/bash
# C4 bench line at N=1 + launch list + DRAM bytes of the main sweep launch
mkdir -p gpurun_out
python bench.py --steps 3 --warmup 3 > gpurun_out/r02g_bench_c4_n1.json 2> gpurun_out/r02g_bench_c4_n1.err; tail -3 gpurun_out/r02g_bench_c4_n1.err; python - <<'PY'
import json
j=json.load(open('gpurun_out/r02g_bench_c4_n1.json'))
print({k:j[k] for k in ('value','ms_per_step','clocks','gpu_launches','parity','cpu_baseline')})
print('e2e',j['e2e']); print('roofline',j['roofline'])
for p in j['graph_pass_roofline']: print(p['pass'],round(p['kernel_ms'],3),'ms',round(p['achieved'],1),'GB/s',round(p['frac'],3))
print('level1',{k:v for k,v in j['level1_p_matrix'].items() if k!='workload'})
PY
