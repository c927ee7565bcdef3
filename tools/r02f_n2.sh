#!/bin/bash
# 2-GPU evidence: the single-process multi-device entry point on two real devices, then the torchrun bench line
timeout 300 python -m pytest tests/test_gpu_multi.py -q -m gpu 2>&1 | tail -2
timeout 600 python tools/multi_probe.py C4 2>&1 | tail -1 | tee gpurun_out/r02f_multi_probe_c4_g2.json
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29632 bench.py --gpus 2 --steps 2 --warmup 3 > gpurun_out/r02f_bench_c4_n2.json 2> gpurun_out/r02f_bench_c4_n2.err
python - <<'PY'
import json
j=json.loads(open("gpurun_out/r02f_bench_c4_n2.json").read().strip().splitlines()[-1])
print(j["value"], j["ms_per_step"], j["e2e"]["value"], j["roofline"]["frac"], j["parity"]["rows_checked"], j["parity"]["id_mismatch_rows"])
PY
