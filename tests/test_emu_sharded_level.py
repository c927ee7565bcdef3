"""multigpu.ShardedLevel with world size 2 on gloo, WITHOUT GPUs: the library is the functional CPU emulation build of the unmodified
kernel sources (tests/emu, "device" memory = host memory), the tensors are CPU tensors, the collectives run on gloo.  Checks the
whole flow of SURVEY 8(e) D1 — broadcast of the walk matrix, slab createSimilarities per rank, the three all-gathers, the level
tail on every rank — against the one-call fused level.  Runs only when the emulation library was built and SPHB_EMULATE=1 is set
(minutes of fiber switching; not part of the default suites):

  python tests/emu/build_emu.py
  SPHB_EMULATE=1 SPHB_LIB=$PWD/tests/emu/_build/libsphb200_emu.so python -m pytest tests/test_emu_sharded_level.py -q
"""
import os
import socket
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
pytestmark = pytest.mark.skipif(os.environ.get("SPHB_EMULATE") != "1" or not os.environ.get("SPHB_LIB"), reason="needs the tests/emu build (SPHB_EMULATE=1, SPHB_LIB)")

WORKER = r'''
import os, sys
import numpy as np
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, os.path.join(sys.argv[1], "tests"))
import torch
import torch.distributed as dist
from sph_b200 import capi
from sph_b200.multigpu import ShardedLevel
from test_gpu_level import walk_like_rows
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
assert capi.load().sphb_version() == -1, "this test drives the emulation build only"
capi.init(0)
n, k, perplexity = 150, 11, 5.0
rp, c, v = walk_like_rows(n, seed=21, hi=30)
class EmuShardedLevel(ShardedLevel):   # test scaffolding: "device" memory of the emulation build is host memory
    def _dev(self):
        return torch.device("cpu")
    def _stream(self, dev):
        return 0
job = EmuShardedLevel(rank, world)
if rank == 0:
    t = (torch.from_numpy(rp), torch.from_numpy(c.view(np.int32)), torch.from_numpy(v))
else:
    t = (None, None, None)
drp, dc, dv = job.replicate_csr(*t, src=0)
assert torch.equal(drp, torch.from_numpy(rp)) and torch.equal(dv, torch.from_numpy(v))
orp, oc, ov = job(drp, dc, dv, k, perplexity, variant=1)
hrp, hc, hv = capi.level_prob_dist(rp, c, v, k, 1e-4, 1, None, True, perplexity, 1, True)
np.testing.assert_array_equal(orp.numpy(), hrp)
np.testing.assert_array_equal(oc.numpy().view(np.uint32), hc)
np.testing.assert_array_equal(ov.numpy(), hv)
dist.barrier()
dist.destroy_process_group()
print("rank", rank, "ok")
'''


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_level_on_gloo_with_the_emulated_kernels(tmp_path, world):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    env = dict(os.environ, OMP_NUM_THREADS="1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1", "--master-port",
           str(_free_port()), str(script), str(ROOT)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=3000, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert r.stdout.count("ok") == world
