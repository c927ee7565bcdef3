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


WORKER_WALKS = r'''
import os, sys
import numpy as np
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, os.path.join(sys.argv[1], "tests"))
import torch.distributed as dist
from sph_b200 import host
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
n = 101
rng = np.random.default_rng(3)
counts = rng.integers(0, 7, n)
counts[40:60] = 0                      # empty rows, also across the slab boundary
rowptr = np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)
col = rng.integers(0, n, rowptr[-1]).astype(np.uint32)
val = rng.uniform(size=rowptr[-1]).astype(np.float32)
full = host.SparseRows(rowptr, col, val)
def walks(m, q0, q1):                  # stand-in for the CUDA slab call: returns the slab of a fixed "result"
    b, e = rowptr[q0], rowptr[q1]
    return rowptr[q0:q1 + 1] - b, col[b:e], val[b:e]
out = host.doRandomWalks(full, host.RandomWalkSettings(), rank=rank, world=world, walks_fn=walks)
np.testing.assert_array_equal(out.rowptr, rowptr)
np.testing.assert_array_equal(out.col, col)
np.testing.assert_array_equal(out.val, val)
dist.barrier()
dist.destroy_process_group()
print("rank", rank, "ok")
'''


WORKER_SIMS = r'''
import os, sys
import numpy as np
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, os.path.join(sys.argv[1], "tests"))
import torch.distributed as dist
import oracle_lib
from sph_b200 import host
from test_gpu_level import walk_like_rows
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
orc = oracle_lib.load()
rp, c, v = walk_like_rows(211, seed=4, lo=3, hi=30, window=40)
m = host.SparseRows(rp, c, v)
k = 9
calls = []
def sims(mat, r0, r1):                 # stand-in for the CUDA slab call: the checker on the same slab contract
    calls.append((r0, r1))
    return orc.create_similarities_rows(mat.rowptr, mat.col, mat.val, k, r0, r1, 1e-4, 1)
cols, vals, nnz = host.createSimilarities(m, k, 1e-4, 1, rank=rank, world=world, sims_fn=sims)
assert calls == [host.shard_rows(211, world)[rank]]
fc, fv, fn = orc.create_similarities(rp, c, v, k, 1e-4, 1)
np.testing.assert_array_equal(nnz, fn)
np.testing.assert_array_equal(cols, fc)
np.testing.assert_array_equal(vals, fv)
assert cols.dtype == np.uint32 and cols.shape == (211, k)
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


def test_sharded_random_walks_world2_gloo(tmp_path):
    import subprocess

    script = tmp_path / "worker_walks.py"
    script.write_text(WORKER_WALKS)
    env = dict(os.environ, OMP_NUM_THREADS="2")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1", "--master-port",
           str(_free_port()), str(script), str(ROOT)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert r.stdout.count("ok") == 2


def test_sharded_create_similarities_world2_gloo(tmp_path):
    """SURVEY 8(e), D1 sharded by output row: every rank computes its slab of rows, one gather, every rank ends with the whole result."""
    import subprocess

    script = tmp_path / "worker_sims.py"
    script.write_text(WORKER_SIMS)
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


@pytest.fixture
def capi_on_oracle(monkeypatch, oracle):
    """Routes the ctypes entry points the host mirror calls to the CPU checker, so that the host-side control flow of
    LevelSimilarities / ImageHierarchy can be exercised here without a GPU (test scaffolding only: the product binds the
    CUDA library and has no such switch)."""
    from sph_b200 import capi

    o = oracle
    monkeypatch.setattr(capi, "level_knn_overlap", lambda ids, ptr, pix, k: o.level_knn_overlap(ids, ptr, pix, k))
    monkeypatch.setattr(capi, "euclid_hausdorff", lambda x, ptr, pix, pa, pb: o.euclid_hausdorff(x, ptr, pix, pa, pb))
    monkeypatch.setattr(capi, "level_knn_walks", lambda rp, c, v, k, mode=0: o.level_knn_walks(rp, c, v, k, mode))
    monkeypatch.setattr(capi, "geodesic_hausdorff",
                        lambda ids, dist, ptr, pix, pa, pb, rowptr=None, labels=None: o.geodesic_hausdorff(rowptr, ids, dist, ptr, pix, pa, pb, labels=labels))
    monkeypatch.setattr(capi, "prow_graph_ragged", lambda rp, ids, dist, perp=0.0, linear=False: o.prow_graph_ragged(rp, ids, dist, perp, linear))

    def level_knn_euclid(x, ptr, pix, k):  # all pairs through the checker's Hausdorff distance, rows ordered by (distance, index)
        n_l = len(ptr) - 1
        cols = np.arange(n_l, dtype=np.int64)
        ids = np.empty((n_l, k), np.int64)
        dist = np.empty((n_l, k), np.float32)
        for r in range(n_l):
            dr = o.euclid_hausdorff(x, ptr, pix, np.full(n_l, r, np.int64), cols)
            order = np.lexsort((cols, dr))[:k]
            ids[r], dist[r] = order, dr[order]
        return ids, dist

    monkeypatch.setattr(capi, "level_knn_euclid", level_knn_euclid)
    monkeypatch.setattr(capi, "rw_as_knn_distances", lambda rp, c, v: o.rw_as_knn_distances(rp, c, v))
    monkeypatch.setattr(capi, "prow_gaussian_flat", lambda d, p: o.prow_gaussian_flat(d, p))
    monkeypatch.setattr(capi, "prow_gaussian_inplace", lambda rp, v, p=0.0, ig=-1: o.prow_gaussian_inplace(rp, v, p, ig))
    monkeypatch.setattr(capi, "remove_zeros", lambda rp, c, v: o.remove_zeros(rp, c, v))
    monkeypatch.setattr(capi, "symmetrize", lambda rp, c, v, umap=False: o.symmetrize(rp, c, v, umap=umap))
    monkeypatch.setattr(capi, "create_similarities", lambda rp, c, v, k, pr=1e-4, variant=0, comp_sizes=None: o.create_similarities(rp, c, v, k, pr, variant, comp_sizes))

    def find_k(rp, c, v, k, top):
        n = len(rp) - 1
        oc = np.full((n, k), 0xFFFFFFFF, np.uint32)
        ov = np.zeros((n, k), np.float32)
        nnz = np.zeros(n, np.int32)
        for i in range(n):
            cc, vv = o.find_k(c[rp[i]:rp[i + 1]], v[rp[i]:rp[i + 1]], k, top)
            oc[i, :len(cc)], ov[i, :len(cc)], nnz[i] = cc, vv, len(cc)
        return oc, ov, nnz

    monkeypatch.setattr(capi, "find_k", find_k)
    return capi


def _toy_hierarchy(oracle):
    import json

    import oracle_lib
    from sph_b200 import host

    g = json.loads((ROOT / "tests" / "golden" / "hierarchy_walks_6x4.json").read_text())
    ids = np.array(g["knn"], np.int64).reshape(24, 4)
    dist = np.array(g["dist"], np.float32).reshape(24, 4)
    graph = host.Graph(ids, dist)
    return host, graph, oracle_lib.OracleOps(oracle)


def test_level_similarities_overlap_control_flow(capi_on_oracle, oracle):
    host, graph, ops = _toy_hierarchy(oracle)
    ih = host.ImageHierarchy(graph, None, 6, 4, ops=ops)
    ih.compute(host.ImageHierarchySettings(componentSim=host.ComponentSim.NEIGH_OVERLAP, usePercentile=False))
    ls = host.LevelSimilarities(graph, hierarchy=ih)
    ls.compute(host.LevelSimilaritiesSettings(levelToCompute=1, componentSim=host.ComponentSim.NEIGH_OVERLAP))
    p = ls.getProbDist(1)
    n1 = ih.numComponents[1]
    assert p.n == n1 and ls.getSimilaritiesGraph(1).knnIndices.shape == (n1, min(ls.getKs()[1], n1))
    dense = np.zeros((n1, n1), np.float32)
    for i in range(n1):
        c, v = p.row(i)
        dense[i, c] = v
    np.testing.assert_array_equal(dense, dense.T)          # symmetrizeTSNE ran
    np.testing.assert_allclose(dense.sum(), n1, rtol=1e-4)  # rows were distributions before the symmetrisation
    assert (p.val != 0).all()                               # zero removal


@pytest.mark.parametrize("kind", ["pair_sims", "topk", "walk_knn", "rw_dist"])
def test_level_similarities_walk_control_flow(capi_on_oracle, oracle, kind):
    host, graph, ops = _toy_hierarchy(oracle)
    ih = host.ImageHierarchy(graph, None, 6, 4, ops=ops)
    ihs = host.ImageHierarchySettings(componentSim=host.ComponentSim.NEIGH_WALKS_SINGLE_OVERLAP, usePercentile=False,
                                      rwHandling=host.RandomWalkHandling.MERGE_RW_NEW_WALKS, normKnnDistances=host.NormalizationScheme.LINEAR)
    ih.compute(ihs, host.RandomWalkSettings(numRandomWalks=10, singleWalkLength=3, parallel=False))
    assert ih.numComponents == [24, 9, 3, 1]
    ls = host.LevelSimilarities(graph, hierarchy=ih)
    lss = host.LevelSimilaritiesSettings(levelToCompute=1, componentSim=host.ComponentSim.NEIGH_WALKS_SINGLE_OVERLAP,
                                         computeSymmetricProbDist=host.NormalizationScheme.NONE)
    if kind == "topk":
        lss.randomWalkPairSims = False
    elif kind == "walk_knn":
        lss.knnFromWalkMetric = True
    elif kind == "rw_dist":
        lss.forceComputeDistances = True
    ls.compute(lss)
    p = ls.getProbDist(1)
    assert p.n == 9 and (np.diff(p.rowptr) >= 0).all() and (p.val != 0).all()
    assert np.isfinite(p.val).all() and (p.val > 0).all()  # (rows need not sum to 1: with fewer entries than the perplexity the
    # reference's search fails and falls back to 1 / (size - 1) per entry, sph/utils/HDILibHelper.hpp:98-104)
    if kind == "walk_knn":
        assert ls.getSimilaritiesGraph(1).knnIndices.shape[0] == 9
    if kind == "rw_dist":
        assert ls.getSimilaritiesGraph(1).numPoints == 9


def test_level_similarities_euclid_control_flow(capi_on_oracle, oracle):
    import oracle_lib
    from sph_b200 import host

    rng = np.random.default_rng(6)
    x = rng.standard_normal((24, 5)).astype(np.float32)
    ids, dist = oracle.knn_reference_graph(x, 4)
    graph = host.Graph(ids, dist)
    ih = host.ImageHierarchy(graph, x, 6, 4, ops=oracle_lib.OracleOps(oracle))
    ih.compute(host.ImageHierarchySettings(componentSim=host.ComponentSim.EUCLID_CENTROID, usePercentile=False))
    assert ih.getNumLevels() >= 3 and ih.numComponents[-1] == 1
    ls = host.LevelSimilarities(graph, hierarchy=ih)
    ls.compute(host.LevelSimilaritiesSettings(levelToCompute=1, componentSim=host.ComponentSim.EUCLID_CENTROID))
    g = ls.getSimilaritiesGraph(1)
    n1 = ih.numComponents[1]
    assert g.knnIndices.shape[0] == n1 and (g.knnIndices[:, 0] == np.arange(n1)).all() and (g.knnDistances[:, 0] == 0).all()
    assert (np.diff(g.knnDistances, axis=1) >= 0).all()
    assert ls.getProbDist(1).n == n1


def _l2_graph(oracle, x, k):
    ids, dist = oracle.knn_reference_graph(x, k)
    return ids, np.sqrt(dist).astype(np.float32)  # L2squared = false: true L2 distances (sph/NearestNeighbors.cpp:224-231)


@pytest.mark.parametrize("sim", ["GEO_CENTROID", "GEO_WALKS"])
def test_image_hierarchy_geodesic_control_flow(capi_on_oracle, oracle, sim):
    """GEO_CENTROID / GEO_WALKS (sph/ImageHierarchy.cpp:192-249 -> componentDistance -> geodesicDistance): the merge loop asks
    for the geodesic distance of every (component, spatial neighbour) pair of a level in one batched call."""
    import oracle_lib
    from sph_b200 import host

    rng = np.random.default_rng(12)
    x = rng.standard_normal((30, 4)).astype(np.float32)
    ids, dist = _l2_graph(oracle, x, 5)
    graph = host.Graph(ids, dist)
    ih = host.ImageHierarchy(graph, x, 6, 5, ops=oracle_lib.OracleOps(oracle))
    ih.compute(host.ImageHierarchySettings(componentSim=getattr(host.ComponentSim, sim), usePercentile=False))
    assert ih.getNumLevels() >= 3 and ih.numComponents[-1] == 1
    # level-0 distances are computeShortestPath values of the 4-neighbour pairs, rows sorted, self first
    d0 = ih.computeDistances(0)
    for c in (0, 7, 29):
        nb = d0.getNeighbors(c)
        dd = d0.getDistances(c)
        assert nb[0] == c and dd[0] == 0 and (np.diff(dd) >= 0).all()
        ref = oracle.shortest_paths(None, ids, dist, np.full(len(nb), c), nb)
        ref[ref < 0] = np.finfo(np.float32).max
        assert np.array_equal(dd, ref)
    # a higher level: the Hausdorff distance over the represented points
    d1 = ih.computeDistances(1)
    ptr, pix = ih.representedDataPoints(1)
    nb = d1.getNeighbors(0)
    ref = oracle.geodesic_hausdorff(None, ids, dist, ptr, pix, np.zeros(len(nb), np.int64), nb)
    ref[nb == 0] = 0
    assert np.array_equal(np.sort(ref), d1.getDistances(0))


def test_image_hierarchy_geodesic_ragged_graph_and_labels(capi_on_oracle, oracle):
    """The GEO_* configurations run on the connected (ragged) graph with the component labels (sph/ComputeHierarchy.cpp:135-144)."""
    import oracle_lib
    from sph_b200 import host

    rng = np.random.default_rng(3)
    x = rng.standard_normal((24, 3)).astype(np.float32)
    x[12:] += 40  # two components
    ids, dist = _l2_graph(oracle, x, 4)
    rp, sid, sdist = oracle.symmetrize_graph(ids, dist)
    graph = host.RaggedGraph(rp, sid, sdist, symmetric=True)
    labels = (np.arange(24) >= 12).astype(np.int64)
    ih = host.ImageHierarchy(graph, x, 6, 4, ops=oracle_lib.OracleOps(oracle))
    ih.compute(host.ImageHierarchySettings(componentSim=host.ComponentSim.GEO_CENTROID, usePercentile=False, componentLabels=labels))
    assert ih.getNumLevels() >= 2
    d0 = ih.computeDistances(0)
    # pixel 11 (row 2, col 3) and pixel 15 (row 3, col 3) are spatial neighbours in different components: no path
    nb, dd = d0.getNeighbors(11), d0.getDistances(11)
    assert dd[list(nb).index(15)] == np.finfo(np.float32).max


def test_level_similarities_geodesic_knn(capi_on_oracle, oracle):
    """LevelSimilarities with GEO_CENTROID: exact level kNN in GeodesicPathSpace, then useKnnDistances (LevelSimilarities.cpp:405-407, :558-561)."""
    import oracle_lib
    from sph_b200 import host

    rng = np.random.default_rng(6)
    x = rng.standard_normal((24, 5)).astype(np.float32)
    ids, dist = _l2_graph(oracle, x, 4)
    graph = host.Graph(ids, dist)
    ih = host.ImageHierarchy(graph, x, 6, 4, ops=oracle_lib.OracleOps(oracle))
    ih.compute(host.ImageHierarchySettings(componentSim=host.ComponentSim.GEO_CENTROID, usePercentile=False))
    ls = host.LevelSimilarities(graph, hierarchy=ih)
    ls.compute(host.LevelSimilaritiesSettings(levelToCompute=1, componentSim=host.ComponentSim.GEO_CENTROID))
    g = ls.getSimilaritiesGraph(1)
    n1 = ih.numComponents[1]
    assert g.knnIndices.shape[0] == n1 and (g.knnIndices[:, 0] == np.arange(n1)).all() and (g.knnDistances[:, 0] == 0).all()
    assert (np.diff(g.knnDistances, axis=1) >= 0).all()
    ptr, pix = ih.representedDataPoints(1)
    row = oracle.geodesic_hausdorff(None, ids, dist, ptr, pix, np.full(n1, 2, np.int64), np.arange(n1))
    order = np.lexsort((np.arange(n1), row))[:g.knnIndices.shape[1]]
    assert np.array_equal(g.knnIndices[2], order) and np.array_equal(g.knnDistances[2], row[order])
    assert ls.getProbDist(1).n == n1


def test_subsample_component_pairs():
    from sph_b200 import host

    ptr = np.array([0, 2, 12, 30], np.int64)
    pix = np.arange(30, dtype=np.int64)
    sptr, spix, spa, spb = host.subsampleComponentPairs(ptr, pix, [0, 1, 2], [1, 2, 2], 5, rng=np.random.default_rng(0))
    assert list(spa[:1]) == [0] and list(spb[:1]) == [1]            # 2 x 10 <= 25: untouched
    assert spa[1] >= 3 and spb[1] >= 3 and spa[2] >= 3 and spb[2] >= 3  # 10 x 18 and 18 x 18: both sides replaced by samples
    for c, src in ((spa[1], 1), (spb[1], 2), (spa[2], 2), (spb[2], 2)):
        lst = spix[sptr[c]:sptr[c + 1]]
        assert len(lst) == 5 and set(lst) <= set(pix[ptr[src]:ptr[src + 1]])
    same = host.subsampleComponentPairs(ptr, pix, [0, 1], [1, 2], None)
    assert same[0] is not None and len(same[0]) == 4
