#!/bin/bash
# streaming sweep (d = 1000, full C5): soft lock-step pacing of the CTAs (profile build: SPHB_KNN_STREAM_PACE = window in 256-row
# tiles, 0 = off; SPHB_KNN_STREAM_STAGGER = start offset per CTA in tiles).  Measured sweep times (ms):
#   pacing off, stagger 4 (the kernel before)  3616      pace 32, stagger 1  3315     pace 16, stagger 1  3267     pace 64, stagger 1  3348
#   pace 32, stagger 0  1989     pace 0, stagger 0  3585     pace 8, stagger 0  1991     pace 16, stagger 0  1968 (default)     pace 64, stagger 0  2126
run() { echo "== $1"; shift; env "$@" SPHB_LIB=$PWD/sph_b200/lib/libsphb200_prof.so PROBE_ITERS=2 timeout 300 python tools/knn_probe.py 1024 1024 1000 91 2>&1 | grep -E "iter 1" | sed -e 's/stats.*accepted\/row/acc\/row/' | cut -c1-300; }
run "pace 0, stagger 0" SPHB_KNN_STREAM_PACE=0 SPHB_KNN_STREAM_STAGGER=0
run "pace 8, stagger 0" SPHB_KNN_STREAM_PACE=8 SPHB_KNN_STREAM_STAGGER=0
run "pace 16, stagger 0" SPHB_KNN_STREAM_PACE=16 SPHB_KNN_STREAM_STAGGER=0
run "pace 64, stagger 0" SPHB_KNN_STREAM_PACE=64 SPHB_KNN_STREAM_STAGGER=0
