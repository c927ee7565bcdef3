#!/bin/bash
# Round-2 evidence run: one bench line per BASELINE.json configuration (C4 is the headline and carries the level-1 block).
mkdir -p gpurun_out
for w in C4 C5 C3 C2 C1; do
  extra="--no-level1"
  [ "$w" = "C4" ] && extra=""
  timeout 900 python bench.py --workload $w --steps 2 --warmup 3 $extra > gpurun_out/r02d_bench_${w,,}_n1.json 2> gpurun_out/r02d_bench_${w,,}_n1.err
  echo "== $w rc=$?"
  python - <<PY
import json
try:
    j = json.loads(open("gpurun_out/r02d_bench_${w,,}_n1.json").read().strip().splitlines()[-1])
    r = j.get("roofline") or {}
    print({k: j.get(k) for k in ("value", "ms_per_step")}, "e2e", (j.get("e2e") or {}).get("value"), "frac", r.get("frac"), "kernel", r.get("kernel"))
    print("parity", j.get("parity"))
    print("cpu", j.get("cpu_baseline"))
except Exception as e:
    print("no line:", e)
PY
  tail -3 gpurun_out/r02d_bench_${w,,}_n1.err
done
