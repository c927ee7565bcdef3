"""SURVEY 8(f)-4 on the device: sphb_geodesic_distances / sphb_geodesic_hausdorff (sph_b200/csrc/geodesic.cu) against the oracle's
restatement of computeShortestPath / geodesicDistance (oracle/oracle_geodesic.cpp).  Bar: bit-exact — the device search and the
oracle's Dijkstra settle the same fixed point (minimum over paths of the left-to-right fp32 sum)."""
import numpy as np
import pytest

from test_oracle_geodesic import knn_graph_l2, two_blobs

pytestmark = pytest.mark.gpu

FMAX = np.finfo(np.float32).max


def random_comps(rng, n, sizes):
    comp_ptr = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
    comp_pix = rng.permutation(n)[:comp_ptr[-1]].astype(np.int64)
    return comp_ptr, comp_pix


def test_point_pairs_match_oracle(capi, oracle):
    x = two_blobs(600, 5, 21)
    ids, dist = knn_graph_l2(x, 8)
    rng = np.random.default_rng(1)
    start = rng.integers(0, 600, 400)
    end = rng.integers(0, 600, 400)
    start[:5] = end[:5]                   # the point itself: slot 0
    end[5:25] = ids[start[5:25], 3]       # stored neighbours: no search
    out = capi.geodesic_distances(ids, dist, start, end)
    ref = oracle.shortest_paths(None, ids, dist, start, end)
    assert np.array_equal(out, ref)
    st = capi.geodesic_last_stats()
    assert st["searches"] > 0 and st["searches"] <= len(np.unique(start)) and st["expanded"] > 0
    assert (out[:5] == 0).all() and np.array_equal(out[5:25], dist[start[5:25], 3])


def test_unreachable_and_labels(capi, oracle):
    x = two_blobs(240, 4, 5, gap=50.0)  # two far blobs: with k = 5 no edge crosses
    ids, dist = knn_graph_l2(x, 5)
    rng = np.random.default_rng(3)
    start = rng.integers(0, 240, 150)
    end = rng.integers(0, 240, 150)
    out = capi.geodesic_distances(ids, dist, start, end)
    ref = oracle.shortest_paths(None, ids, dist, start, end)
    assert np.array_equal(out, ref)
    assert (out == -1).sum() > 30
    labels = (np.arange(240) >= 120).astype(np.int64)
    out_l = capi.geodesic_distances(ids, dist, start, end, labels=labels)
    assert np.array_equal(out_l, oracle.shortest_paths(None, ids, dist, start, end, labels=labels))
    # labels cut pairs the graph itself would connect
    labels2 = (np.arange(240) % 2).astype(np.int64)
    out2 = capi.geodesic_distances(ids, dist, start, end, labels=labels2)
    assert np.array_equal(out2, oracle.shortest_paths(None, ids, dist, start, end, labels=labels2))
    assert (out2 == -1).sum() > (out == -1).sum()


def test_hausdorff_matches_oracle(capi, oracle):
    x = two_blobs(500, 6, 8)
    ids, dist = knn_graph_l2(x, 9)
    rng = np.random.default_rng(5)
    sizes = [1, 1, 3, 0, 7, 12, 20, 33, 40, 2]
    comp_ptr, comp_pix = random_comps(rng, 500, sizes)
    comp_pix[comp_ptr[5] + 1] = comp_pix[comp_ptr[5]]  # a point listed twice (sampling with replacement)
    pa = np.array([0, 0, 2, 4, 8, 6, 5, 3, 7, 7, 7, 9, 1], np.int64)
    pb = np.array([1, 0, 4, 7, 6, 6, 1, 2, 8, 5, 2, 3, 9], np.int64)
    out = capi.geodesic_hausdorff(ids, dist, comp_ptr, comp_pix, pa, pb)
    ref = oracle.geodesic_hausdorff(None, ids, dist, comp_ptr, comp_pix, pa, pb)
    assert np.array_equal(out, ref)
    assert out[1] == 0 and out[7] == 0 and out[11] == 0  # a component with itself; empty components
    assert capi.geodesic_last_stats()["sources"] == sum(sizes[c] for c in np.unique(pa))


def test_hausdorff_ragged_graph_with_labels(capi, oracle):
    x = two_blobs(300, 4, 13, gap=30.0)
    ids, dist = knn_graph_l2(x, 6)
    # a ragged graph: every point keeps its row and receives its reverse edges (what symmetrizeGraph's output looks like)
    rows = [[(int(ids[i, j]), float(dist[i, j])) for j in range(6)] for i in range(300)]
    for i in range(300):
        for j in range(1, 6):
            t = int(ids[i, j])
            if all(v != i for v, _ in rows[t]):
                rows[t].append((i, float(dist[i, j])))
    rows[7].append((299, np.float32(1.5)))  # one bridge between the blobs, one way
    rowptr = np.concatenate([[0], np.cumsum([len(r) for r in rows])]).astype(np.int64)
    rid = np.array([v for r in rows for v, _ in r], np.int64)
    rdist = np.array([w for r in rows for _, w in r], np.float32)
    labels = (np.arange(300) >= 150).astype(np.int64)
    rng = np.random.default_rng(9)
    comp_ptr, comp_pix = random_comps(rng, 300, [5, 9, 14, 1, 25, 30])
    pa = np.repeat(np.arange(6), 6).astype(np.int64)
    pb = np.tile(np.arange(6), 6).astype(np.int64)
    for lab in (None, labels):
        out = capi.geodesic_hausdorff(rid, rdist, comp_ptr, comp_pix, pa, pb, rowptr=rowptr, labels=lab)
        ref = oracle.geodesic_hausdorff(rowptr, rid, rdist, comp_ptr, comp_pix, pa, pb, labels=lab)
        assert np.array_equal(out, ref)
    assert (ref == FMAX).any()


def test_level0_grid_pairs(capi, oracle):
    """The level-0 call of ImageHierarchy::computeDistances: every pixel against itself and its 4 spatial neighbours."""
    h, w = 20, 24
    rng = np.random.default_rng(17)
    x = rng.normal(size=(h * w, 6)).astype(np.float32)
    ids, dist = knn_graph_l2(x, 7)
    pa, pb = [], []
    for r in range(h):
        for c in range(w):
            p = r * w + c
            pa.append(p)
            pb.append(p)
            for dr, dc in ((-1, 0), (0, 1), (1, 0), (0, -1)):
                if 0 <= r + dr < h and 0 <= c + dc < w:
                    pa.append(p)
                    pb.append((r + dr) * w + c + dc)
    n = h * w
    comp_ptr = np.arange(n + 1, dtype=np.int64)
    comp_pix = np.arange(n, dtype=np.int64)
    out = capi.geodesic_hausdorff(ids, dist, comp_ptr, comp_pix, pa, pb)
    ref = oracle.geodesic_hausdorff(None, ids, dist, comp_ptr, comp_pix, np.array(pa), np.array(pb))
    assert np.array_equal(out, ref)
    raw = capi.geodesic_distances(ids, dist, pa, pb)
    assert np.array_equal(np.where(raw < 0, FMAX, raw), out)


def test_zero_weight_edges_and_duplicates(capi, oracle):
    rng = np.random.default_rng(4)
    x = rng.normal(size=(200, 3)).astype(np.float32)
    x[50:60] = x[40:50]  # exact duplicates: zero-weight edges
    ids, dist = knn_graph_l2(x, 6)
    start = rng.integers(0, 200, 120)
    end = rng.integers(0, 200, 120)
    assert np.array_equal(capi.geodesic_distances(ids, dist, start, end), oracle.shortest_paths(None, ids, dist, start, end))
    # all weights zero: the bucket width degenerates, the search must still end
    z = np.zeros_like(dist)
    assert np.array_equal(capi.geodesic_distances(ids, z, start, end), oracle.shortest_paths(None, ids, z, start, end))


def test_larger_graph(capi, oracle):
    rng = np.random.default_rng(23)
    x = rng.normal(size=(3000, 8)).astype(np.float32)
    ids, dist = knn_graph_l2(x, 12)
    start = rng.integers(0, 3000, 256)
    end = rng.integers(0, 3000, 256)
    out = capi.geodesic_distances(ids, dist, start, end)
    assert np.array_equal(out, oracle.shortest_paths(None, ids, dist, start, end))
    l2 = np.sqrt(((x[start] - x[end]) ** 2).sum(-1))
    ok = out >= 0
    assert ok.sum() > 200 and (out[ok] - l2[ok] >= -1e-4).all()  # tests/AStarTest.cpp:160


def test_resident_graph_handle_equals_one_shot_calls(capi, oracle):
    """sphb_geodesic_graph_create: the graph is uploaded once and serves several calls (every level of a hierarchy)."""
    x = two_blobs(500, 6, 8)
    ids, dist = knn_graph_l2(x, 9)
    labels = (np.arange(500) % 7 != 0).astype(np.int64)
    rng = np.random.default_rng(11)
    comp_ptr, comp_pix = random_comps(rng, 500, [4, 9, 1, 16, 25])
    pa = np.array([0, 1, 2, 3, 4, 4], np.int64)
    pb = np.array([1, 3, 0, 4, 0, 4], np.int64)
    start, end = rng.integers(0, 500, 100), rng.integers(0, 500, 100)
    for lab in (None, labels):
        with capi.GeodesicGraph(ids, dist, labels=lab) as g:
            for _ in range(2):  # the handle survives a call
                assert np.array_equal(g.distances(start, end), capi.geodesic_distances(ids, dist, start, end, labels=lab))
                assert np.array_equal(g.hausdorff(comp_ptr, comp_pix, pa, pb), capi.geodesic_hausdorff(ids, dist, comp_ptr, comp_pix, pa, pb, labels=lab))
        assert np.array_equal(capi.geodesic_distances(ids, dist, start, end, labels=lab), oracle.shortest_paths(None, ids, dist, start, end, labels=lab))


@pytest.mark.parametrize("sim", ["GEO_CENTROID", "GEO_WALKS"])
def test_image_hierarchy_geodesic_device_equals_oracle_ops(capi, oracle, sim):
    """The host merge loop with the geodesic metric (ImageHierarchy::computeDistances -> geodesicDistance, sph/ImageHierarchy.cpp:192-249)
    on the device ops (resident graph handle, one batched call per level) builds the hierarchy the oracle-backed ops build."""
    import oracle_lib
    from sph_b200 import host

    rng = np.random.default_rng(31)
    x = rng.normal(size=(8 * 6, 5)).astype(np.float32)
    ids, dist = knn_graph_l2(x, 6)
    graph = host.Graph(ids, dist)
    settings = host.ImageHierarchySettings(componentSim=getattr(host.ComponentSim, sim), usePercentile=False)
    rws = host.RandomWalkSettings(numRandomWalks=8, singleWalkLength=3)
    a = host.ImageHierarchy(graph, x, 8, 6)
    a.compute(settings, rws)
    b = host.ImageHierarchy(graph, x, 8, 6, ops=oracle_lib.OracleOps(oracle))
    b.compute(settings, rws)
    assert a.getNumLevels() == b.getNumLevels() >= 3 and a.numComponents == b.numComponents
    for la, lb in zip(a.parents, b.parents):
        assert np.array_equal(la, lb)
    d_a, d_b = a.computeDistances(1), b.computeDistances(1)
    assert np.array_equal(d_a.knnIndices, d_b.knnIndices) and np.array_equal(d_a.knnDistances, d_b.knnDistances)


def test_bad_arguments(capi):
    from sph_b200.capi import SphbError

    ids = np.array([[0, 1], [1, 0]], np.int64)
    with pytest.raises(SphbError):
        capi.geodesic_distances(ids, np.array([[0, -1], [0, 1]], np.float32), [0], [1])
    with pytest.raises(SphbError):
        capi.geodesic_distances(ids, np.array([[0, 1], [0, 1]], np.float32), [0], [2])
    assert len(capi.geodesic_distances(ids, np.array([[0, 1], [0, 1]], np.float32), [], [])) == 0
    # nothing but empty components: no search, no column minima, distance 0 as utils::euclidDistance's batched form returns it
    out = capi.geodesic_hausdorff(ids, np.array([[0, 1], [0, 1]], np.float32), [0, 0, 0], np.zeros(0, np.int64), [0, 1], [1, 1])
    assert out.tolist() == [0, 0]
