"""the oracle's geodesic distances (oracle/oracle_geodesic.cpp).  Boost.Graph is absent here, so the pins are
the reference's own test bar (tests/AStarTest.cpp:160-164: its A* and Boost's agree to 1e-4 relative, and a geodesic is never
shorter than the L2 distance) applied to the restated A*, and scipy's Dijkstra as an independent implementation."""
import numpy as np
import pytest


def knn_graph_l2(x, k):
    """Exact kNN graph with TRUE L2 distances (L2squared = false, the reference default), self first."""
    x64 = x.astype(np.float64)
    sq = (x64 * x64).sum(1)
    d2 = np.maximum(sq[:, None] + sq[None, :] - 2.0 * (x64 @ x64.T), 0.0)
    d2[np.arange(len(x)), np.arange(len(x))] = 0.0
    ids = np.argsort(d2, axis=1, kind="stable")[:, :k]
    for i in range(len(x)):  # ensureClosestPointIsSelf for duplicate points
        if ids[i, 0] != i:
            j = np.where(ids[i] == i)[0]
            ids[i, j[0] if len(j) else k - 1] = ids[i, 0]
            ids[i, 0] = i
    dist = np.sqrt(np.take_along_axis(d2, ids, 1)).astype(np.float32)
    return ids.astype(np.int64), dist


def two_blobs(n, d, seed, gap=0.0):
    rng = np.random.default_rng(seed)
    x = rng.normal(size=(n, d)).astype(np.float32)
    x[n // 2:] += np.float32(gap)
    return x


@pytest.fixture(scope="module")
def small():
    x = two_blobs(400, 5, 3)
    ids, dist = knn_graph_l2(x, 8)
    return x, ids, dist


def scipy_all(ids, dist, sources):
    from scipy.sparse import csr_matrix
    from scipy.sparse.csgraph import dijkstra

    n, k = ids.shape
    rows = np.repeat(np.arange(n), k)
    keep = ids.reshape(-1) != rows
    w = dist.reshape(-1).astype(np.float64)
    # csr_matrix sums duplicate entries and drops explicit zeros: keep the smallest of parallel edges, lift exact zeros
    order = np.lexsort((w[keep], ids.reshape(-1)[keep], rows[keep]))
    r, c, v = rows[keep][order], ids.reshape(-1)[keep][order], w[keep][order]
    first = np.ones(len(r), bool)
    first[1:] = (r[1:] != r[:-1]) | (c[1:] != c[:-1])
    g = csr_matrix((np.maximum(v[first], 1e-300), (r[first], c[first])), shape=(n, n))
    return dijkstra(g, directed=True, indices=sources)


def test_sssp_matches_scipy(oracle, small):
    _, ids, dist = small
    src = [0, 17, 399]
    ref = scipy_all(ids, dist, src)
    for s, row in zip(src, ref):
        got = oracle.sssp(None, ids, dist, s).astype(np.float64)
        reach = np.isfinite(row)
        assert (got[~reach] == np.finfo(np.float32).max).all()
        np.testing.assert_allclose(got[reach], row[reach], rtol=2e-6)  # float sums along <= ~20 hops against fp64


def test_shortest_paths_shortcuts(oracle, small):
    _, ids, dist = small
    n, k = ids.shape
    # self: slot 0 of the own row; direct neighbour: the stored distance of the FIRST matching slot
    start = np.array([5, 5, 9], np.int64)
    end = np.array([5, ids[5, 3], ids[9, k - 1]], np.int64)
    out = oracle.shortest_paths(None, ids, dist, start, end)
    assert out[0] == 0 and out[1] == dist[5, 3] and out[2] == dist[9, k - 1]
    # labels: different components give -1 without a search
    labels = np.zeros(n, np.int64)
    labels[n // 2:] = 1
    out = oracle.shortest_paths(None, ids, dist, [0], [n - 1], labels=labels)
    assert out[0] == -1


def test_unreachable_is_minus_one(oracle):
    x = two_blobs(120, 4, 5, gap=50.0)  # two far blobs: with k = 5 no edge crosses
    ids, dist = knn_graph_l2(x, 5)
    out = oracle.shortest_paths(None, ids, dist, [0, 1], [119, 2])
    assert out[0] == -1 and out[1] > 0


def test_astar_agrees_with_dijkstra_to_the_reference_bar(oracle, small):
    x, ids, dist = small
    rng = np.random.default_rng(8)
    start = rng.integers(0, len(x), 200)
    end = rng.integers(0, len(x), 200)
    a = oracle.astar(ids, dist, x, start, end)
    dj = oracle.shortest_paths(None, ids, dist, start, end)
    found = (a >= 0) & (dj >= 0)
    assert (a < 0).sum() == (dj < 0).sum() == (~found).sum()
    assert found.sum() > 150
    np.testing.assert_allclose(a[found], dj[found], rtol=1e-4)  # tests/AStarTest.cpp:163
    l2 = np.sqrt(((x[start] - x[end]) ** 2).sum(-1))
    assert (a[found] - l2[found] >= -1e-4).all()  # tests/AStarTest.cpp:160


def test_geodesic_hausdorff_matches_definition(oracle, small):
    _, ids, dist = small
    n = len(ids)
    rng = np.random.default_rng(2)
    perm = rng.permutation(n)[:120]
    sizes = [1, 1, 3, 7, 12, 20, 30, 46]
    comp_ptr = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
    comp_pix = perm[:comp_ptr[-1]].astype(np.int64)
    pa = np.array([0, 0, 2, 3, 7, 5, 4], np.int64)
    pb = np.array([1, 0, 3, 6, 5, 5, 1], np.int64)
    out = oracle.geodesic_hausdorff(None, ids, dist, comp_ptr, comp_pix, pa, pb)
    for p in range(len(pa)):
        a = comp_pix[comp_ptr[pa[p]]:comp_ptr[pa[p] + 1]]
        b = comp_pix[comp_ptr[pb[p]]:comp_ptr[pb[p] + 1]]
        m = np.array([oracle.shortest_paths(None, ids, dist, np.full(len(b), s), b) for s in a])
        m[m < 0] = np.finfo(np.float32).max
        assert out[p] == max(m.min(1).max(), m.min(0).max())


def test_sssp_matches_the_committed_scipy_distances(oracle):
    """tests/golden/geodesic_scipy_60.json (make_external_goldens.py): an independent Dijkstra, pinned without scipy at test time."""
    import json
    from pathlib import Path

    g = json.loads((Path(__file__).parent / "golden" / "geodesic_scipy_60.json").read_text())
    ids = np.array(g["ids"], np.int64)
    dist = np.array([[float.fromhex(v) for v in r] for r in g["dist_f32_hex"]], np.float32)
    for s, row in zip(g["sources"], g["distances"]):
        got = oracle.sssp(None, ids, dist, s).astype(np.float64)
        for t, want in enumerate(row):
            if want is None:
                assert got[t] == np.finfo(np.float32).max
            else:
                assert abs(got[t] - want) <= 2e-6 * max(want, 1.0)
