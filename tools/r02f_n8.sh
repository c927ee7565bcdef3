#!/bin/bash
# 8-GPU evidence: torchrun bench line, then the single-process multi-device entry point on all eight devices
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29633 bench.py --gpus 8 --steps 2 --warmup 3 > gpurun_out/r02f_bench_c4_n8.json 2> gpurun_out/r02f_bench_c4_n8.err
python - <<'PY'
import json
j=json.loads(open("gpurun_out/r02f_bench_c4_n8.json").read().strip().splitlines()[-1])
print(j["value"], j["ms_per_step"], j["e2e"]["value"], j["roofline"]["frac"], j["parity"]["rows_checked"], j["parity"]["id_mismatch_rows"], j["clocks"])
PY
timeout 300 python tools/multi_probe.py C4 2>&1 | tail -1 | tee gpurun_out/r02f_multi_probe_c4_g8.json
