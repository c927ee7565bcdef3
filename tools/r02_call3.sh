#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_knn.py tests/test_gpu_prow.py tests/test_gpu_level.py -m gpu -x -q 2>&1 | tail -8
run() { echo "== $1"; shift; env "$@" PROBE_ITERS=3 timeout 300 python tools/knn_probe.py 2048 2048 128 91 303104 2>&1 | grep iter | sed -e 's/stats.*accepted\/row/acc\/row/' | cut -c1-260; }
run "bisect (default)" X=1
run "sort-select (round 1)" SPHB_LIB=$PWD/sph_b200/lib/libsphb200_sortsel.so
SPHB_LIB=$PWD/sph_b200/lib/libsphb200_prof.so SPHB_KNN_PROF=1 PROBE_ITERS=1 timeout 300 python tools/knn_probe.py 2048 2048 128 91 303104 2>&1 | grep -E "knn prof" | tee gpurun_out/r02_sweep_prof_bisect.txt
timeout 120 python tools/prow_probe.py 1048576 91 2>&1 | tail -2
timeout 120 python tools/prow_probe.py 1048576 91 3.0 2>&1 | tail -1
PROBE_REPS=2 timeout 300 python tools/level1_probe.py > gpurun_out/r02_level1_probe_after2.json 2> gpurun_out/r02_level1_probe_after2.err; python - <<'PY'
import json
j=json.load(open('gpurun_out/r02_level1_probe_after2.json'))
print({k:v for k,v in j.items() if 'rep1' in k})
PY
