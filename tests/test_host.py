"""CPU tests (no GPU) of the host-side mirror: settings/schedule logic against the oracle, row sharding,
and the world_size-2 gloo path of the sharded kNN (compute injected, since there is no GPU here)."""
import os
import socket
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent


def test_shard_rows_cover_and_balance():
    from sph_b200 import host

    for n, w in [(10, 3), (4194304, 8), (7, 8), (0, 2), (21025, 4)]:
        slabs = host.shard_rows(n, w)
        assert slabs[0][0] == 0 and slabs[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(slabs[:-1], slabs[1:]))
        sizes = [b - a for a, b in slabs]
        assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        host.shard_rows(5, 0)


def test_level_schedule_matches_oracle(oracle):
    from sph_b200 import host

    for sizes, k0 in [([21025, 5000, 900, 40], 91), ([4194304, 1200000, 300000, 70000, 9000, 500, 12], 91), ([111104, 30000], 301), ([500, 100], 16)]:
        perps, ks = host.level_schedule(sizes, k0)
        operps, oks = oracle.level_schedule(sizes, k0)
        np.testing.assert_allclose(perps, operps, rtol=1e-7)
        assert ks == list(oks)


def test_num_neighbors_from_setting():
    from sph_b200 import host

    # evaluation/RunEvaluation.cpp:241-248
    assert host.numNeighborsFromSetting(21025, 90) == 91
    assert host.numNeighborsFromSetting(21025, 0) == 301
    assert host.numNeighborsFromSetting(500, 0) == 31
    assert host.NearestNeighbors.indexHeuristic(5000) == host.KnnIndex.BruteForce
    assert host.NearestNeighbors.indexHeuristic(50000) == host.KnnIndex.IVFFlat


def test_renorm_l2_matches_oracle(oracle):
    from sph_b200 import host

    rng = np.random.default_rng(0)
    x = rng.standard_normal((50, 17)).astype(np.float32)
    x[3] = 0
    out = np.zeros_like(x)
    import ctypes as C

    oracle.lib.orc_renorm_l2(x.ctypes.data_as(C.POINTER(C.c_float)), 50, 17, out.ctypes.data_as(C.POINTER(C.c_float)))
    np.testing.assert_allclose(host.renorm_l2(x), out, rtol=2e-7, atol=0)


def test_nearest_neighbors_error_convention(caplog):
    from sph_b200 import host

    nn = host.NearestNeighbors(np.zeros((10, 3), np.float32))
    nn.compute(host.NearestNeighborsSettings(numNearestNeighbors=0))
    assert not nn.isValid() and nn.getKnnGraph() is None  # log + early return, outputs untouched

    def boom(*a):
        raise RuntimeError("no device")

    nn = host.NearestNeighbors(np.zeros((10, 3), np.float32), knn_fn=boom)
    nn.compute(host.NearestNeighborsSettings(numNearestNeighbors=3))
    assert not nn.isValid() and nn.getKnnGraph() is None


def test_symmetric_neighbors_path(oracle, monkeypatch):
    # NearestNeighbors::compute with symmetricNeighbors = true runs computeSymmetrizedNnGraph right after the search
    # (sph/NearestNeighbors.cpp:176-177); the compute calls are injected with the checker (no GPU here)
    from sph_b200 import host

    rng = np.random.default_rng(2)
    x = rng.standard_normal((120, 5)).astype(np.float32)

    def knn(x, k, metric, l2sq, q0, q1):
        return oracle.knn_reference_graph(x, k, l2_squared=l2sq, rows=np.arange(q0, q1), metric=metric)

    nn = host.NearestNeighbors(x, knn_fn=knn)
    monkeypatch.setattr(nn, "_symmetrize", lambda ids, dist: oracle.symmetrize_graph(ids, dist))
    nn.compute(host.NearestNeighborsSettings(numNearestNeighbors=6, symmetricNeighbors=True))
    g = nn.getSymGraph()
    assert g is not None and g.symmetric and g.numPoints == 120
    assert (g.nns >= 6).all()
    for p in range(120):
        assert g.getNeighbors(p)[0] == p
        for q in g.getNeighbors(p):
            assert p in g.getNeighbors(int(q))
    nn2 = host.NearestNeighbors(x, knn_fn=knn)
    nn2.compute(host.NearestNeighborsSettings(numNearestNeighbors=6))
    assert nn2.getSymGraph() is None


def test_scatter_flat_rows():
    # sph/utils/HDILibHelper.hpp:131-139: slot 0 skipped, -1 skipped, later duplicate overwrites, rows column-sorted
    from sph_b200 import host

    ids = np.array([[0, 5, 2, -1], [1, 3, 3, 0]], np.int64)
    temp = np.array([[9, .5, .25, .7], [9, .1, .2, .3]], np.float32)
    rp, c, v = host.scatter_flat_rows(ids, temp)
    assert rp.tolist() == [0, 2, 4] and c.tolist() == [2, 5, 0, 3]
    np.testing.assert_array_equal(v, np.array([.25, .5, .3, .2], np.float32))


WORKER = r'''
import os, sys
import numpy as np
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, os.path.join(sys.argv[1], "tests"))
import torch.distributed as dist
import oracle_lib
from sph_b200 import host
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
orc = oracle_lib.load()
rng = np.random.default_rng(5)
x = rng.standard_normal((301, 6)).astype(np.float32)
calls = []
def knn(x, k, metric, l2sq, q0, q1):   # stand-in for the CUDA slab call: the checker on the same slab contract
    calls.append((q0, q1))
    return orc.knn_reference_graph(x, k, l2_squared=l2sq, rows=np.arange(q0, q1), metric=metric)
nn = host.NearestNeighbors(x, knn_fn=knn)
nn.compute(host.NearestNeighborsSettings(numNearestNeighbors=7), rank=rank, world=world)
assert nn.isValid()
assert calls == [host.shard_rows(301, world)[rank]]
full_ids, full_dist = orc.knn_reference_graph(x, 7)
np.testing.assert_array_equal(nn.getKnnIndices(), full_ids)
np.testing.assert_array_equal(nn.getKnnDistances(), full_dist)
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


def test_sharded_knn_world2_gloo(tmp_path):
    import subprocess

    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    env = dict(os.environ, OMP_NUM_THREADS="2")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1", "--master-port",
           str(_free_port()), str(script), str(ROOT)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert r.stdout.count("ok") == 2


def test_bench_stdout_guard(tmp_path):
    # bench.py keeps its stdout to ONE JSON line: while NCCL creates its communicator, fd 1 points at stderr
    code = (
        "import ctypes, sys, bench\n"
        "with bench.stdout_to_stderr():\n"
        "    ctypes.CDLL(None).puts(b'banner from C')\n"
        "    print('banner from python')\n"
        "print('{\"ok\": 1}')\n"
    )
    import subprocess
    import sys

    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, cwd=str(ROOT), timeout=120)
    assert r.returncode == 0, r.stderr
    assert r.stdout.strip() == '{"ok": 1}'
    assert "banner from C" in r.stderr and "banner from python" in r.stderr
