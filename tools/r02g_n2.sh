#!/bin/bash
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29634 bench.py --gpus 2 --steps 2 --warmup 3 > gpurun_out/r02g_bench_c4_n2.json 2> gpurun_out/r02g_bench_c4_n2.err
python - <<'PY'
import json
j=json.loads(open("gpurun_out/r02g_bench_c4_n2.json").read().strip().splitlines()[-1])
print(j["value"], j["ms_per_step"], j["e2e"]["value"], j["e2e"]["ms_per_step"], j["roofline"]["frac"], j["parity"]["rows_checked"], j["parity"]["id_mismatch_rows"], j["parity"]["e2e_equals_resident_first_rows"])
PY
tail -2 gpurun_out/r02g_bench_c4_n2.err
