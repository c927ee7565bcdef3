#!/bin/bash
# What to run first on hardware in the next round: everything written after the GPU budget of round 2 was
# spent, in the order of its risk.  One gpurun call per block; nothing here is a measurement under a profiler.
#
#   1 GPU:   gpurun --timeout 400 -- 'bash tools/r03_first_calls.sh one'
#   2 GPUs:  gpurun --gpus 2 --timeout 400 -- 'bash tools/r03_first_calls.sh two'
#   8 GPUs:  gpurun --gpus 8 --timeout 900 -- 'bash tools/r03_first_calls.sh eight'
set -u
mkdir -p gpurun_out
case "${1:-one}" in
one)
  # the row-range createSimilarities kernels, the level tail, the resident geodesic graph, the filtered relaxations
  timeout 200 python -m pytest tests/test_zx_gpu_geodesic.py tests/test_zy_gpu_sims_slab.py tests/test_gpu_level.py tests/test_gpu_graph.py -q -m gpu \
      -p no:cacheprovider 2>&1 | tee gpurun_out/r03_new_tests.log | tail -5
  timeout 60 python tools/geodesic_probe.py --level1 --cpu-pairs 20000 --out gpurun_out/r03_geodesic_probe.jsonl 2>&1 | tail -6   # 454 ms before the filters
  timeout 120 python tools/level1_sharded_probe.py --workload C2 --out gpurun_out/r03_level1_sharded_g1.json 2>&1 | tail -2
  ;;
two)
  timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 \
      tools/level1_sharded_probe.py --workload C2 --out gpurun_out/r03_level1_sharded_g2.json 2>&1 | tail -3
  timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29542 \
      tools/level1_c4_sharded.py --shape 512,216,204,16,1002 --band-rows 64 --out gpurun_out/r03_level1_c2_sharded_g2.json 2>&1 | tail -3
  ;;
eight)
  # the north_star target: C4 kNN + level-1 similarities + P matrix on 8 GPUs (one GPU: 117 s, tools/level1_c4.py)
  timeout 800 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29543 \
      tools/level1_c4_sharded.py --workload C4 --out gpurun_out/r03_level1_c4_sharded_g8.json 2>&1 | tail -3
  ;;
esac
