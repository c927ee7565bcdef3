"""CPU tests of the host half of NearestNeighbors::connectComponents (sph/NearestNeighbors.cpp:494-861) with the
device calls replaced by the checker: the reference's own test (tests/UtilsTests.cpp:662-703: S-curve, 1000 points, k = 3 ->
more than one component before, exactly one after), the spanning tree against scipy, and insertDistance's rules."""
import numpy as np
import pytest

from sph_b200 import host


def s_curve(n, noise, seed):
    """utils::createSCurve (sph/utils/TestData.cpp:88-140) in numpy: same construction, numpy's generator (the reference fills it
    inside an OpenMP loop from one mt19937_64, so its point order is not reproducible either)."""
    rng = np.random.default_rng(seed)
    t = 3.0 * np.pi * (rng.random(n) - 0.5)
    x = np.sin(t)
    y = np.sign(t) * (np.cos(t) - 1.0)
    z = 2.0 * rng.random(n)
    p = np.stack([x, y, z], axis=1) + noise * rng.standard_normal((n, 3))
    return p.astype(np.float32)


def _nn_on_oracle(oracle, x, k, symmetric=False, monkeypatch=None):
    def knn(x, k, metric, l2sq, q0, q1):
        return oracle.knn_reference_graph(x, k, l2_squared=l2sq, rows=np.arange(q0, q1), metric=metric)

    nn = host.NearestNeighbors(x, knn_fn=knn)
    if symmetric:
        monkeypatch.setattr(nn, "_symmetrize", lambda ids, dist: oracle.symmetrize_graph(ids, dist))
    nn.compute(host.NearestNeighborsSettings(numNearestNeighbors=k, symmetricNeighbors=symmetric))
    if symmetric:
        g = nn.getSymGraph()
        nn._cc = oracle.weak_components(g.knnIndices, g.rowptr)
    else:
        nn._cc = oracle.weak_components(nn.getKnnGraph().knnIndices)
    return nn


@pytest.mark.parametrize("symmetric", [False, True])
def test_connecting_components_knn_graph(oracle, monkeypatch, symmetric):
    x = s_curve(1000, 0.2, 123)
    nn = _nn_on_oracle(oracle, x, 3, symmetric, monkeypatch)
    num, labels = nn.getConnectedComponents()
    assert num > 1
    g = nn.connectComponents(bridge_ops=oracle)
    assert g is not None and g.numPoints == 1000
    new_num, _ = oracle.weak_components(g.knnIndices, g.rowptr)
    assert new_num == 1
    # exactly one bidirectional edge per tree edge, unless the edge existed already in one direction
    base_nnz = len(nn.getSymGraph().knnIndices) if symmetric else 1000 * 3
    added = len(g.knnIndices) - base_nnz
    assert num - 1 <= added <= 2 * (num - 1)
    for p in range(1000):
        nb, ds = g.getNeighbors(p), g.getDistances(p)
        assert nb[0] == p and ds[0] == 0 and len(set(nb.tolist())) == len(nb)
        assert (np.diff(ds) >= 0).all()     # checkDistancesAreNonDecreasing (:833)


def test_single_component_graph_is_copied_unchanged(oracle):
    rng = np.random.default_rng(0)
    x = rng.standard_normal((200, 4)).astype(np.float32)
    nn = _nn_on_oracle(oracle, x, 12)
    assert nn.getConnectedComponents()[0] == 1
    g = nn.connectComponents(bridge_ops=oracle)
    np.testing.assert_array_equal(g.knnIndices, nn.getKnnGraph().knnIndices.reshape(-1))
    np.testing.assert_array_equal(g.knnDistances, nn.getKnnGraph().knnDistances.reshape(-1))


def test_spanning_tree_matches_scipy():
    from scipy.sparse.csgraph import minimum_spanning_tree

    rng = np.random.default_rng(5)
    c = rng.standard_normal((300, 7)).astype(np.float32)
    edges = host.centroid_spanning_tree(c)
    assert edges.shape == (299, 2) and (edges[:, 0] < edges[:, 1]).all()
    w = np.sqrt(((c[edges[:, 0]] - c[edges[:, 1]]) ** 2).sum(axis=1))
    assert (np.diff(w) >= -1e-6).all()                        # Kruskal emits the edges in ascending weight
    full = np.sqrt(((c[:, None, :].astype(np.float64) - c[None, :, :]) ** 2).sum(axis=2))
    ref = minimum_spanning_tree(full).tocoo()
    ref_edges = {(min(a, b), max(a, b)) for a, b in zip(ref.row.tolist(), ref.col.tolist())}
    assert ref_edges == {tuple(e) for e in edges.tolist()}
    assert host.centroid_spanning_tree(c[:1]).shape == (0, 2)


def test_insert_distance_rules():
    # 3 rows of a fixed-k graph (self first)
    rowptr = np.array([0, 3, 6, 9], np.int64)
    ids = np.array([0, 1, 2, 1, 0, 2, 2, 1, 0], np.int64)
    dist = np.array([0, 1, 2, 0, 1, 3, 0, 0, 0], np.float32)
    ia = np.array([0, 0, 2, 1], np.int64)
    ib = np.array([5, 1, 7, 1], np.int64)
    dd = np.array([1.0, 0.5, 0.0, 9.0], np.float32)
    rp, ni, nd = host.insert_graph_edges(rowptr, ids, dist, ia, ib, dd)
    # (0,5,1.0): behind the last distance <= 1 -> [0,1,5,2]; (5,0): row 5 does not exist in this toy graph -> guarded by the caller's ids
    # (0,1,0.5): 1 is already a neighbour of 0 -> skipped, and (1,0) likewise; (2,7,0.0): upper_bound = end of the zero run, never slot 0
    # (1,1): a == b -> skipped
    assert ni[rp[0]:rp[1]].tolist() == [0, 1, 5, 2] and nd[rp[0]:rp[1]].tolist() == [0, 1, 1, 2]
    assert ni[rp[1]:rp[2]].tolist() == [1, 0, 2]
    assert ni[rp[2]:rp[3]].tolist() == [2, 1, 0, 7] and nd[rp[2]:rp[3]].tolist() == [0, 0, 0, 0]


def test_bridge_oracle_centres_and_pairs(oracle):
    rng = np.random.default_rng(9)
    x = rng.standard_normal((90, 5)).astype(np.float32)
    labels = rng.integers(0, 4, size=90).astype(np.int64)
    c = oracle.component_centers(x, labels, 4)
    for k in range(4):
        np.testing.assert_allclose(c[k], x[labels == k].astype(np.float64).mean(axis=0), rtol=1e-6)
    ia, ib, dd = oracle.bridge_components(x, labels, 4, np.array([[0, 1], [2, 3]]))
    for e, (ca, cb) in enumerate([(0, 1), (2, 3)]):
        A, B = np.nonzero(labels == ca)[0], np.nonzero(labels == cb)[0]
        full = ((x[B][:, None, :].astype(np.float64) - x[A][None, :, :]) ** 2).sum(axis=2)
        bi, ai = np.unravel_index(np.argmin(full), full.shape)
        assert (ia[e], ib[e]) == (A[ai], B[bi])
        np.testing.assert_allclose(dd[e], full[bi, ai], rtol=1e-6)
