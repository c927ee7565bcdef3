#!/bin/bash
# Evidence for the code written after the GPU budget of round 2 was spent: everything below runs the UNMODIFIED kernel sources
# under the CPU emulation of tests/emu (logic check, never a measurement).  Output: profiles/r02l_emulation_runs.log
set -u
cd "$(dirname "$0")/.."
LOG=profiles/r02l_emulation_runs.log
python tests/emu/build_emu.py > /dev/null || exit 1
export SPHB_EMULATE=1 SPHB_LIB=$PWD/tests/emu/_build/libsphb200_emu.so OMP_NUM_THREADS=1
{
echo "== git $(git rev-parse --short HEAD), $(date -u +%FT%TZ), emulation library: tests/emu/_build/libsphb200_emu.so (sphb_version() == -1)"
echo "== tests/test_zx_gpu_geodesic.py tests/test_zy_gpu_sims_slab.py (the -m gpu tests, kernels emulated)"
python -m pytest tests/test_zx_gpu_geodesic.py tests/test_zy_gpu_sims_slab.py -m gpu -v -p no:cacheprovider 2>&1 | grep -E "PASSED|FAILED|SKIPPED|ERROR|passed|failed"
echo "== tests/test_emu_sharded_level.py (multigpu.ShardedLevel, world 2 and 3, gloo)"
python -m pytest tests/test_emu_sharded_level.py -v -p no:cacheprovider 2>&1 | grep -E "PASSED|FAILED|SKIPPED|ERROR|passed|failed"
for G in 1 2 3; do
  echo "== tools/level1_c4_sharded.py --emulate, $G rank(s)"
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port $((29530+G)) tools/level1_c4_sharded.py \
      --emulate --shape 12,16,6,3,5 --k 7 --band-rows 4 --walks 6,3 --parity-rows 16 2>/dev/null | grep '^{'
done
echo "== tests/emu/geo_race_model.cpp (8 OS threads as the warps of a CTA vs sequential Dijkstra)"
g++ -O2 -std=c++17 -pthread tests/emu/geo_race_model.cpp -o /tmp/geo_race_model && /tmp/geo_race_model 40
} > "$LOG" 2>&1
tail -5 "$LOG"
