#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_knn.py -m gpu -x -q 2>&1 | tail -3
python bench.py --steps 3 --warmup 3 > gpurun_out/r02b_bench_c4_n1.json 2> gpurun_out/r02b_bench_c4_n1.err; tail -3 gpurun_out/r02b_bench_c4_n1.err; python - <<'PY'
import json
j=json.load(open('gpurun_out/r02b_bench_c4_n1.json'))
print({k:j[k] for k in ('value','ms_per_step','parity')})
print('e2e',j['e2e']['value']); print('roofline',{k:j['roofline'][k] for k in ('frac','kernel_ms','rerank_ms','fallback_ms','prep_ms')})
PY
