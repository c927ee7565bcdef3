"""GPU parity tests of the ragged-graph passes (csrc/graph_merge.cu) through the C ABI vs the CPU oracle and the
reference's golden vectors: symmetrizeGraph (F1), mergeGraphNodes / mergeNodesDataDistances (RW3),
useRandomWalksAsKnnDistances (D9), zero removal (B7), exact level kNN in the random-walk metrics (D7),
Euclid-Hausdorff component distance (D6).  All integer / index results bit-exact; fp32 values bit-exact too (every value
is either copied or produced by the same fp32 / fp64 operation chain as the oracle's).

STATUS: hardware-verified.  All 42 tests passed on a B200 in the round-1 driver run (GPUTEST_r01.json, reported as XPASS
because the module then still carried xfail(strict=False)); the mark is gone, a regression now fails the GPU gate.
"""
import json
from pathlib import Path

import numpy as np
import pytest

from oracle_lib import dense_to_csr

pytestmark = [pytest.mark.gpu]
GOLD = Path(__file__).resolve().parent / "golden"


def gold(name):
    return json.loads((GOLD / name).read_text())


def eq3(a, b):
    for x, y in zip(a, b):
        np.testing.assert_array_equal(x, y)


def random_knn_graph(n, k, seed, dup_frac=0.0):
    """Fixed-k graph with self first, distances ascending; `dup_frac` of the points are exact duplicates of another
    point (distance 0 to it), the situation ensureClosestPointIsSelf exists for."""
    rng = np.random.default_rng(seed)
    ids = np.empty((n, k), np.int64)
    dist = np.empty((n, k), np.float32)
    for i in range(n):
        nb = rng.choice(np.delete(np.arange(n), i), k - 1, replace=False)
        d = np.sort(rng.integers(1, 40, k - 1)).astype(np.float32) / 8  # many ties within and across rows
        ids[i] = np.concatenate([[i], nb])
        dist[i] = np.concatenate([[0], d])
    for i in rng.choice(n, int(dup_frac * n), replace=False):
        dist[i, 1] = 0.0  # a zero-distance neighbour with possibly smaller index than i
    return ids, dist


def random_stochastic(n, nnz_per_row, seed, diag=True):
    rng = np.random.default_rng(seed)
    rowptr = [0]
    col, val = [], []
    for i in range(n):
        m = int(rng.integers(0, nnz_per_row + 1))
        pool = np.arange(n) if diag else np.delete(np.arange(n), i)
        c = np.sort(rng.choice(pool, size=min(m, len(pool)), replace=False))
        w = rng.uniform(0.05, 1.0, len(c))
        col.extend(c.tolist())
        val.extend((w / max(w.sum(), 1e-9)).tolist())
        rowptr.append(len(col))
    return np.array(rowptr, np.int64), np.array(col, np.uint32), np.array(val, np.float32)


# ---------------------------------------------------------------- F1
def test_symmetrize_graph_golden(capi):
    g = gold("symmetrize_graph_24.json")
    ids = np.array(g["ids"], np.int64).reshape(-1, g["k"])
    dist = np.array(g["dist"], np.float32).reshape(-1, g["k"])
    rp, oi, od = capi.symmetrize_graph(ids, dist)
    for p, exp in g["rows"].items():
        p = int(p)
        assert oi[rp[p]:rp[p + 1]].tolist() == exp


@pytest.mark.parametrize("n,k,dup", [(24, 3, 0.0), (2000, 16, 0.05), (5000, 91, 0.01)])
def test_symmetrize_graph_matches_oracle(capi, oracle, n, k, dup):
    ids, dist = random_knn_graph(n, k, seed=n + k, dup_frac=dup)
    eq3(capi.symmetrize_graph(ids, dist), oracle.symmetrize_graph(ids, dist))


def test_symmetrize_graph_ragged_and_empty_rows(capi, oracle):
    # ragged input (a symmetrised graph symmetrised again is a fixed point), rows without entries, self missing
    ids, dist = random_knn_graph(300, 8, seed=2)
    rp, oi, od = oracle.symmetrize_graph(ids, dist)
    eq3(capi.symmetrize_graph(oi, od, rp), (rp, oi, od))
    rp2 = np.array([0, 2, 2, 5, 6], np.int64)
    ids2 = np.array([1, 2, 0, 2, 3, 3], np.int64)      # row 0 lacks itself (shift branch), row 1 is empty
    dist2 = np.array([1, 2, 0.5, 0, 1, 0], np.float32)
    eq3(capi.symmetrize_graph(ids2, dist2, rp2), oracle.symmetrize_graph(ids2, dist2, rp2))


# ---------------------------------------------------------------- RW3
def test_merge_graph_nodes_golden(capi):
    g = gold("merge_graph_9to3.json")
    ids = np.array(g["ids"], np.int64).reshape(-1, g["k"])
    dist = np.array(g["dist"], np.float32).reshape(-1, g["k"])
    rp, oi, od = capi.merge_graph_nodes(ids, dist, g["parents"], 3)
    for m, exp in g["expected"].items():
        m = int(m)
        assert oi[rp[m]] == m and od[rp[m]] == 0
        for t, (nid, nd) in enumerate(exp, start=1):
            assert oi[rp[m] + t] == nid
            np.testing.assert_allclose(od[rp[m] + t], nd, rtol=1e-6)


@pytest.mark.parametrize("n,k,nm", [(9, 4, 3), (3000, 12, 700), (20000, 31, 9000)])
def test_merge_graph_nodes_matches_oracle(capi, oracle, n, k, nm):
    ids, dist = random_knn_graph(n, k, seed=n)
    rng = np.random.default_rng(n + 1)
    parents = rng.integers(0, nm, n).astype(np.uint64)
    parents[:nm] = np.arange(nm)  # every merged node has a child
    eq3(capi.merge_graph_nodes(ids, dist, parents, nm), oracle.merge_graph_nodes(ids, dist, parents, nm))


def test_merge_data_distances_matches_oracle(capi, oracle):
    rp, c, v = random_stochastic(2500, 20, seed=8)
    v[::13] = 0.0  # stored zeros are "empty" for the reference's coeff() test
    rng = np.random.default_rng(3)
    parents = rng.integers(0, 600, 2500).astype(np.uint64)
    parents[:600] = np.arange(600)
    eq3(capi.merge_data_distances(rp, c, v, parents, 600), oracle.merge_data_distances(rp, c, v, parents, 600))


# ---------------------------------------------------------------- D9 / B7
def test_rw_as_knn_distances_matches_oracle(capi, oracle):
    rp, c, v = random_stochastic(4000, 40, seed=12)
    v[::11] = 0.0
    v = (np.round(v * 16) / 16).astype(np.float32)  # heavy ties: the stable order is what is tested
    eq3(capi.rw_as_knn_distances(rp, c, v), oracle.rw_as_knn_distances(rp, c, v))


def test_remove_zeros_matches_oracle(capi, oracle):
    rp, c, v = random_stochastic(4000, 40, seed=13)
    v[::3] = 0.0
    eq3(capi.remove_zeros(rp, c, v), oracle.remove_zeros(rp, c, v))
    z = np.zeros_like(v)
    orp, oc, ov = capi.remove_zeros(rp, c, z)
    assert orp[-1] == 0 and len(oc) == 0 and (orp == 0).all()


# ---------------------------------------------------------------- D7
@pytest.mark.parametrize("mode", [0, 1])
@pytest.mark.parametrize("n,per_row,k", [(50, 5, 12), (1500, 40, 31), (600, 3, 100)])
def test_level_knn_walks_matches_oracle(capi, oracle, mode, n, per_row, k):
    rp, c, v = random_stochastic(n, per_row, seed=n + mode)
    ids, dist = capi.level_knn_walks(rp, c, v, k, mode)
    rids, rdist = oracle.level_knn_walks(rp, c, v, k, mode)
    np.testing.assert_array_equal(ids, rids)
    np.testing.assert_array_equal(dist, rdist)  # same fp32 summation order as the per-pair loop


def test_level_knn_walks_k_larger_than_n(capi, oracle):
    rp, c, v = random_stochastic(20, 6, seed=1)
    ids, dist = capi.level_knn_walks(rp, c, v, 33, 0)
    rids, rdist = oracle.level_knn_walks(rp, c, v, 33, 0)
    np.testing.assert_array_equal(ids, rids)
    np.testing.assert_array_equal(dist, rdist)
    assert (ids[:, 20:] == -1).all()


# ---------------------------------------------------------------- D6
@pytest.mark.parametrize("d", [3, 7, 128, 200])
def test_euclid_hausdorff_matches_oracle(capi, oracle, d):
    rng = np.random.default_rng(d)
    n = 3000
    x = rng.normal(size=(n, d)).astype(np.float32)
    sizes = rng.integers(0, 60, 120)
    sizes[0] = 1
    sizes[1] = 1
    comp_ptr = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
    comp_pix = rng.integers(0, n, comp_ptr[-1]).astype(np.int64)
    pa = rng.integers(0, 120, 500).astype(np.int64)
    pb = rng.integers(0, 120, 500).astype(np.int64)
    pa[:2], pb[:2] = [0, 0], [1, 0]  # level-0 style single-pixel pairs, one of them identical
    out = capi.euclid_hausdorff(x, comp_ptr, comp_pix, pa, pb)
    ref = oracle.euclid_hausdorff(x, comp_ptr, comp_pix, pa, pb)
    np.testing.assert_array_equal(out, ref)
    assert out[1] == 0


@pytest.mark.parametrize("d,ncomp,k", [(3, 200, 31), (64, 150, 150), (16, 5000, 12)])
def test_level_knn_euclid_matches_all_pairs(capi, oracle, d, ncomp, k):
    """D7 in the Euclid metric: every component against every component, k smallest (distance, index) per row — against the
    oracle's Hausdorff distance over all pairs (rows sampled where ncomp^2 is large)."""
    rng = np.random.default_rng(ncomp + d)
    n = 4 * ncomp
    x = rng.normal(size=(n, d)).astype(np.float32)
    x[n // 2:] = x[:n - n // 2]            # duplicated pixels: zero distances and ties
    sizes = rng.integers(1, 6, ncomp)
    comp_ptr = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
    comp_pix = rng.integers(0, n, comp_ptr[-1]).astype(np.int64)
    ids, dist = capi.level_knn_euclid(x, comp_ptr, comp_pix, k)
    cols = np.arange(ncomp, dtype=np.int64)
    rows = np.arange(ncomp) if ncomp <= 200 else rng.choice(ncomp, 40, replace=False)
    for r in rows:
        dr = oracle.euclid_hausdorff(x, comp_ptr, comp_pix, np.full(ncomp, r, np.int64), cols)
        order = np.lexsort((cols, dr))[:k]
        np.testing.assert_array_equal(ids[r], order)
        np.testing.assert_array_equal(dist[r], dr[order])
    assert (np.diff(dist, axis=1) >= 0).all()


# ---------------------------------------------------------------- host mirror on top of the new entry points
def test_level_similarities_walk_knn_path(capi, oracle):
    """LevelSimilarities with rwHandling = MERGE_RW_NEW_WALKS_AND_KNN: exact level kNN in the Bhattacharyya metric, then
    useKnnDistances (flat Gaussian rows, tol 1e-5) and zero removal (sph/LevelSimilarities.cpp:211-252, 510-515, 566-581)."""
    from sph_b200 import host

    n0, k0 = 600, 31
    ids, dist = random_knn_graph(n0, k0, seed=4)
    rw1 = host.SparseRows(*random_stochastic(150, 25, seed=6))
    ls = host.LevelSimilarities(host.Graph(ids, dist), [None, rw1])
    lss = host.LevelSimilaritiesSettings(levelToCompute=1, knnFromWalkMetric=True, componentSim=host.ComponentSim.NEIGH_WALKS,
                                         computeSymmetricProbDist=host.NormalizationScheme.NONE)
    ls.compute(lss)
    g = ls.getSimilaritiesGraph(1)
    k1 = ls.getKs()[1]
    rids, rdist = oracle.level_knn_walks(rw1.rowptr, rw1.col, rw1.val, k1, 0)
    np.testing.assert_array_equal(g.knnIndices, rids)
    np.testing.assert_array_equal(g.knnDistances, rdist)
    temp = oracle.prow_gaussian_flat(rdist, ls.getPerplexities()[1])
    rp, c, v = host.scatter_flat_rows(rids, temp)
    zrp, zc, zv = oracle.remove_zeros(rp, c, v)
    p = ls.getProbDist(1)
    np.testing.assert_array_equal(p.rowptr, zrp)
    np.testing.assert_array_equal(p.col, zc)
    np.testing.assert_allclose(p.val, zv, rtol=1e-5, atol=1e-12)


def test_level_similarities_topk_transition_path(capi, oracle):
    """randomWalkPairSims = false: the k_l largest transition values per row, then the in-place Gaussian rows
    (sph/LevelSimilarities.cpp:481-497)."""
    from sph_b200 import host

    ids, dist = random_knn_graph(600, 31, seed=5)
    rw1 = host.SparseRows(*random_stochastic(150, 60, seed=7))
    ls = host.LevelSimilarities(host.Graph(ids, dist), [None, rw1])
    ls.compute(host.LevelSimilaritiesSettings(levelToCompute=1, randomWalkPairSims=False, computeSymmetricProbDist=host.NormalizationScheme.NONE))
    k1, perp1 = ls.getKs()[1], ls.getPerplexities()[1]
    rows_c, rows_v, rp = [], [], [0]
    for i in range(rw1.n):
        c, v = rw1.row(i)
        oc, ov = oracle.find_k(c, v, k1, True)
        rows_c.append(oc)
        rows_v.append(ov)
        rp.append(rp[-1] + len(oc))
    rp = np.array(rp, np.int64)
    c = np.concatenate(rows_c).astype(np.uint32)
    v = oracle.prow_gaussian_inplace(rp, np.concatenate(rows_v).astype(np.float32), perp1, -1)
    zrp, zc, zv = oracle.remove_zeros(rp, c, v)
    p = ls.getProbDist(1)
    np.testing.assert_array_equal(p.rowptr, zrp)
    np.testing.assert_array_equal(p.col, zc)
    np.testing.assert_allclose(p.val, zv, rtol=1e-5, atol=1e-12)


# ---------------------------------------------------------------- D7, neighbour-overlap metric (the default componentSim)
def random_components(n, ncomp, seed):
    rng = np.random.default_rng(seed)
    lab = rng.integers(0, ncomp, n)
    lab[:ncomp] = np.arange(ncomp)
    order = np.argsort(lab, kind="stable")
    ptr = np.concatenate([[0], np.cumsum(np.bincount(lab, minlength=ncomp))]).astype(np.int64)
    return ptr, order.astype(np.int64)


@pytest.mark.parametrize("n,kd,ncomp,k", [(24, 3, 9, 5), (3000, 12, 400, 31), (2000, 8, 1500, 64), (300, 6, 20, 33)])
def test_level_knn_overlap_matches_oracle(capi, oracle, n, kd, ncomp, k):
    ids, _ = random_knn_graph(n, kd, seed=n + kd)
    ptr, pix = random_components(n, ncomp, seed=ncomp)
    a = capi.level_knn_overlap(ids, ptr, pix, k)
    b = oracle.level_knn_overlap(ids, ptr, pix, k)
    np.testing.assert_array_equal(a[0], b[0])
    np.testing.assert_array_equal(a[1], b[1])  # integer counts, one fp32 division and one subtraction: bit-exact


def test_level_similarities_default_overlap_path(capi, oracle):
    """The default configuration (componentSim = NEIGH_OVERLAP on both stages): hierarchy on the device, level-1 kNN in the
    overlap metric, useKnnDistances (flat Gaussian rows, tol 1e-5), zero removal, symmetrizeTSNE — vs the oracle chain."""
    import oracle_lib
    from sph_b200 import host, synth

    x = synth.hyperspectral(20, 16, 12, endmembers=4, seed=8)
    ids, dist = oracle.knn_reference_graph(x, 9)  # input only: the device kNN is bit-identical to it (test_gpu_knn.py)
    g = host.Graph(ids, dist)
    ih = host.ImageHierarchy(g, x, 20, 16)
    ih.compute(host.ImageHierarchySettings(componentSim=host.ComponentSim.NEIGH_OVERLAP, usePercentile=False))
    assert ih.getNumLevels() >= 3
    ls = host.LevelSimilarities(g, hierarchy=ih)
    ls.compute(host.LevelSimilaritiesSettings(levelToCompute=1, componentSim=host.ComponentSim.NEIGH_OVERLAP))
    ptr, pix = ih.representedDataPoints(1)
    k1, perp1 = min(ls.getKs()[1], len(ptr) - 1), ls.getPerplexities()[1]
    rids, rdist = oracle.level_knn_overlap(ids, ptr, pix, k1)
    sg = ls.getSimilaritiesGraph(1)
    np.testing.assert_array_equal(sg.knnIndices, rids)
    np.testing.assert_array_equal(sg.knnDistances, rdist)
    temp = oracle.prow_gaussian_flat(rdist, perp1)
    rp, c, v = host.scatter_flat_rows(rids, temp)
    rp, c, v = oracle.remove_zeros(rp, c, v)
    orp, oc, ov = oracle.symmetrize(rp, c, v)
    p = ls.getProbDist(1)
    np.testing.assert_array_equal(p.rowptr, orp)
    np.testing.assert_array_equal(p.col, oc)
    np.testing.assert_allclose(p.val, ov, rtol=1e-5, atol=1e-12)


# ---------------------------------------------------------------- F2 connected components
def test_weak_components_golden(capi):
    g = gold("components_9.json")
    num, labels = capi.weak_components(np.array(g["ids"], np.int64).reshape(-1, g["k"]))
    assert num == g["num_components"] and labels.tolist() == g["labels"]


@pytest.mark.parametrize("n,k,clusters", [(24, 3, 1), (3000, 4, 37), (20000, 6, 500), (5000, 2, 5000)])
def test_weak_components_matches_oracle(capi, oracle, n, k, clusters):
    # neighbours drawn inside `clusters` index-interleaved groups (long chains included): many components of uneven size
    rng = np.random.default_rng(n + k)
    group = np.arange(n) % clusters
    ids = np.empty((n, k), np.int64)
    for i in range(n):
        same = np.nonzero(group == group[i])[0]
        lo = max(0, np.searchsorted(same, i) - 2)
        ids[i] = np.concatenate([[i], rng.choice(same[lo:lo + 5], k - 1)])  # near neighbours in the group: long thin components
    num, labels = capi.weak_components(ids)
    rnum, rlabels = oracle.weak_components(ids)
    assert num == rnum
    np.testing.assert_array_equal(labels, rlabels)


def test_connected_components_of_symmetric_graph(capi, oracle):
    from sph_b200 import host

    ids, dist = random_knn_graph(1500, 3, seed=77)
    ids[:, 1:] = (np.arange(1500)[:, None] // 50) * 50 + (ids[:, 1:] % 50)  # neighbours inside the row's block of 50: >= 30 components
    nn = host.NearestNeighbors(np.zeros((1500, 2), np.float32), knn_fn=lambda *a: (ids, dist))
    nn.compute(host.NearestNeighborsSettings(numNearestNeighbors=3, symmetricNeighbors=True, computeConnectComponents=True))
    num, labels = nn.getConnectedComponents()
    g = nn.getSymGraph()
    rnum, rlabels = oracle.weak_components(g.knnIndices, g.rowptr)
    assert num == rnum and num > 1
    np.testing.assert_array_equal(labels, rlabels)


# ---------------------------------------------------------------- B2 / B5 on ragged graphs
@pytest.mark.parametrize("linear", [False, True])
def test_prow_graph_ragged_matches_oracle(capi, oracle, linear):
    # the ragged graph a symmetrisation produces (rows of 8 .. ~40 neighbours), plus single-entry and two-entry rows
    ids, dist = random_knn_graph(1200, 8, seed=21)
    rp, oi, od = oracle.symmetrize_graph(ids, dist)
    extra_ids = np.array([1200, 1201, 0], np.int64)
    extra_dist = np.array([0, 0, 2.5], np.float32)
    rp = np.concatenate([rp, [rp[-1] + 1, rp[-1] + 3]]).astype(np.int64)
    oi = np.concatenate([oi, extra_ids])
    od = np.concatenate([od, extra_dist])
    a = capi.prow_graph_ragged(rp, oi, od, 0.0, linear)
    b = oracle.prow_graph_ragged(rp, oi, od, 0.0, linear)
    np.testing.assert_array_equal(a[0], b[0])
    np.testing.assert_array_equal(a[1], b[1])
    np.testing.assert_allclose(a[2], b[2], rtol=1e-5, atol=1e-12)


def test_hierarchy_merge_data_new_walks_on_device(capi):
    from sph_b200 import host

    g = gold("hierarchy_walks_6x4.json")
    ids = np.array(g["knn"], np.int64).reshape(24, 4)
    dist = np.array(g["dist"], np.float32).reshape(24, 4)
    ih = host.ImageHierarchy(host.Graph(ids, dist), None, 6, 4)
    ihs = host.ImageHierarchySettings(componentSim=host.ComponentSim.NEIGH_WALKS, usePercentile=False,
                                      rwHandling=host.RandomWalkHandling.MERGE_DATA_NEW_WALKS, normKnnDistances=host.NormalizationScheme.LINEAR)
    ih.compute(ihs, host.RandomWalkSettings(numRandomWalks=10, singleWalkLength=3))
    assert ih.getNumLevels() >= 2 and ih.numComponents[-1] == 1
    for lvl, mg in enumerate(ih.mergedDataGraphs):
        assert mg.numPoints == ih.numComponents[lvl]
        assert (mg.knnIndices[mg.rowptr[:-1]] == np.arange(mg.numPoints)).all()
    for rw in ih.randomWalks:
        np.testing.assert_allclose(np.add.reduceat(rw.val, rw.rowptr[:-1]), 1.0, atol=1e-3)


# ---------------------------------------------------------------- RW1 sharded by start points
def test_random_walk_slabs_concatenate_to_the_full_result(capi):
    rp, c, v = random_stochastic(700, 12, seed=31, diag=False)
    full = capi.random_walks(rp, c, v, 20, 6, capi.W_LINEAR, seed=5)
    parts = [capi.random_walks_slab(rp, c, v, 20, 6, q0, q1, capi.W_LINEAR, seed=5) for q0, q1 in [(0, 233), (233, 234), (234, 700)]]
    rowptr = np.concatenate([[0]] + [p[0][1:] + off for p, off in zip(parts, np.cumsum([0] + [len(p[1]) for p in parts[:-1]]))])
    np.testing.assert_array_equal(rowptr, full[0])
    np.testing.assert_array_equal(np.concatenate([p[1] for p in parts]), full[1])
    np.testing.assert_array_equal(np.concatenate([p[2] for p in parts]), full[2])  # Philox keyed by the global start id
    e = capi.random_walks_slab(rp, c, v, 20, 6, 100, 100, capi.W_LINEAR, seed=5)
    assert e[0].tolist() == [0] and len(e[1]) == 0


def test_symmetrize_graph_dev_resident(capi, oracle):
    """Device-pointer variant on buffers that stay on the device (torch tensors on a B200; plain host arrays under tests/emu,
    where device memory is host memory)."""
    import os

    n, k = 3000, 12
    ids, dist = random_knn_graph(n, k, seed=5, dup_frac=0.02)
    cap = 2 * n * k
    ref = oracle.symmetrize_graph(ids, dist)
    if os.environ.get("SPHB_EMULATE") == "1":
        orp, oi, od = np.zeros(n + 1, np.int64), np.zeros(cap, np.int64), np.zeros(cap, np.float32)
        m = capi.symmetrize_graph_dev(n, 0, k, ids.ctypes.data, dist.ctypes.data, n * k, orp.ctypes.data, oi.ctypes.data, od.ctypes.data, cap, 0)
        got = (orp, oi[:m], od[:m])
    else:
        import torch

        capi.init(0)
        d_ids, d_dist = torch.from_numpy(ids).cuda(), torch.from_numpy(dist).cuda()
        orp = torch.zeros(n + 1, dtype=torch.int64, device="cuda")
        oi = torch.zeros(cap, dtype=torch.int64, device="cuda")
        od = torch.zeros(cap, dtype=torch.float32, device="cuda")
        torch.cuda.synchronize()
        m = capi.symmetrize_graph_dev(n, 0, k, d_ids.data_ptr(), d_dist.data_ptr(), n * k, orp.data_ptr(), oi.data_ptr(), od.data_ptr(), cap,
                                      capi.stream_ptr())
        torch.cuda.synchronize()
        got = (orp.cpu().numpy(), oi[:m].cpu().numpy(), od[:m].cpu().numpy())
    eq3(got, ref)


def test_merge_graph_nodes_long_row_path(capi, oracle):
    # one merged node absorbs half of the graph: its row (> 2048 distinct neighbours) takes the global-sort path, the rest the
    # shared-memory path is exercised by the other tests (rows <= 2048)
    n = 6000
    ids, dist = random_knn_graph(n, 12, seed=41)
    parents = np.concatenate([np.zeros(3000), np.arange(1, 3001)]).astype(np.uint64)
    a = capi.merge_graph_nodes(ids, dist, parents, 3001)
    b = oracle.merge_graph_nodes(ids, dist, parents, 3001)
    assert np.diff(b[0]).max() > 2048
    eq3(a, b)
