#!/bin/bash
# probe the CTA stagger parameter on the one-wave C4-like slab
for s in 0 1 4 16 64; do
  echo "== stagger $s"; SPHB_KNN_STAGGER=$s PROBE_ITERS=2 python tools/knn_probe.py 1024 1024 128 91 37888 2>&1 | tail -1
done
