#!/bin/bash
# resident sweep (C4): stagger / pacing window after the streaming kernel's result (profile build).  Slab of 303 104 rows, sweep ms:
#   stagger 4, pace 256 (default until then) 313.2 | stagger 0, pace 256: 309.6 | stagger 0, pace 64: 303.5 | stagger 1, pace 64: 303.5 |
#   stagger 0, pace 32: 307.9 | stagger 0, pace 16: 308.7.  Full C4: stagger 4 / pace 256: 4428 | stagger 0 / pace 64: 4326 (adopted) | stagger 0 / pace 32: 4336
run() { echo "== $1"; shift; env "$@" SPHB_LIB=$PWD/sph_b200/lib/libsphb200_prof.so PROBE_ITERS=2 timeout 300 python tools/knn_probe.py 2048 2048 128 91 ${ROWS:-303104} 2>&1 | grep -E "iter 1" | sed -e 's/stats.*accepted\/row/acc\/row/' | cut -c1-260; }
run "stagger 0, pace 32" SPHB_KNN_STAGGER=0 SPHB_KNN_PACE=32
run "stagger 0, pace 16" SPHB_KNN_STAGGER=0 SPHB_KNN_PACE=16
ROWS=4194304
run "full C4: stagger 4, pace 256" X=1
run "full C4: stagger 0, pace 64" SPHB_KNN_STAGGER=0 SPHB_KNN_PACE=64
run "full C4: stagger 0, pace 32" SPHB_KNN_STAGGER=0 SPHB_KNN_PACE=32
