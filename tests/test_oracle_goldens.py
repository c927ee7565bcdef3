"""CPU tests (no GPU): the oracle against the reference's own golden vectors.

Each test re-expresses one Catch2 test of alxvth/SPH (file:line in tests/golden/README.md)
on the CPU restatement under oracle/.  These pin the checker; the `-m gpu` tests then
compare the CUDA path with the checker.
"""
import json
from pathlib import Path

import numpy as np
import pytest

import oracle_lib
from oracle_lib import csr_to_dense, dense_to_csr

GOLD = Path(__file__).resolve().parent / "golden"


def gold(name):
    return json.loads((GOLD / name).read_text())


# ---------------------------------------------------------------- Stage B6
@pytest.mark.parametrize("umap", [False, True])
def test_symmetrize_golden(oracle, umap):
    g = gold("symmetrize_5x5.json")
    rp, c, v = dense_to_csr(g["input"])
    orp, oc, ov = oracle.symmetrize(rp, c, v, umap=umap)
    out = csr_to_dense(orp, oc, ov)
    exp = np.array(g["umap" if umap else "tsne"], np.float32)
    np.testing.assert_allclose(out, exp, atol=g["tol"])
    np.testing.assert_array_equal(out, out.T)  # isSymmetric
    # column-sorted rows (MapMemEff order)
    for i in range(5):
        assert (np.diff(oc[orp[i]:orp[i + 1]].astype(np.int64)) > 0).all()


# ---------------------------------------------------------------- Stage C (RW2)
def test_merge_nodes_synth_golden(oracle):
    g = gold("merge_nodes_synth.json")
    rp, c, v = dense_to_csr(g["input"])
    orp, oc, ov = oracle.rw_merge(rp, c, v, g["parents"], 3, norm=g["norm"], weight_by_size=g["weight_by_size"])
    np.testing.assert_array_equal(csr_to_dense(orp, oc, ov), np.array(g["expected"], np.float32))


def test_merge_weight_by_size_rows_sum_to_one(oracle):
    # mergeNodesRandomWalks(weightBySize) divides by the summed weights; with norm the rows are L1-normalised
    rng = np.random.default_rng(3)
    m = (rng.uniform(size=(40, 40)) < 0.2) * rng.uniform(0.1, 1, size=(40, 40))
    m /= np.maximum(m.sum(1, keepdims=True), 1e-9)
    rp, c, v = dense_to_csr(m)
    parents = rng.integers(0, 7, size=40)
    parents[:7] = np.arange(7)
    orp, oc, ov = oracle.rw_merge(rp, c, v, parents, 7, norm=True, weight_by_size=True)
    d = csr_to_dense(orp, oc, ov, 7)
    np.testing.assert_allclose(d.sum(1), 1.0, atol=1e-5)


# ---------------------------------------------------------------- Stage D
def test_similarities_5x5_three_way(oracle):
    g = gold("sims_5x5.json")
    a = np.array(g["input"], np.float32)
    for r in a:
        assert r.sum(dtype=np.float32) == np.float32(1.0)  # REQUIRE(row.sum() == 1.0f)
    rp, c, v = dense_to_csr(a)
    n = 5
    # createSimilaritiesSPH (explicit sqrt dots) == -log(sqrt(A) sqrt(A)^T) computed directly
    srp, sc, sv = oracle.create_similarities_sph(rp, c, v, prune_val=g["prune"])
    sph = csr_to_dense(srp, sc, sv)
    s = np.sqrt(a)
    with np.errstate(divide="ignore"):
        direct = -np.log((s @ s.T).astype(np.float32))
    mask = (s @ s.T) > 0
    np.fill_diagonal(mask, False)
    np.testing.assert_allclose(sph[mask], direct[mask], rtol=1e-4, atol=1e-5)
    # Eigen path + bottom-k conversion must be bit-equal to the HDI path (UtilsTests.cpp:1012-1027)
    k = g["k"]
    ec, ev, en = oracle.create_similarities(rp, c, v, k, prune_val=g["prune"], variant=2)
    hc, hv, hn = oracle.create_similarities(rp, c, v, k, prune_val=g["prune"], variant=1)
    np.testing.assert_array_equal(en, hn)
    for i in range(n):
        np.testing.assert_array_equal(ec[i, :en[i]], hc[i, :hn[i]])
        np.testing.assert_array_equal(ev[i, :en[i]], hv[i, :hn[i]])
    # and the kept entries are the k smallest distances of the explicit matrix, L1-normalised
    for i in range(n):
        cand = [(sph[i, j], j) for j in range(n) if j != i and mask[i, j]]
        cand.sort()
        keep = sorted(j for _, j in cand[:k])
        assert list(hc[i, :hn[i]]) == keep
        vals = np.array([sph[i, j] for j in keep], np.float64)
        np.testing.assert_allclose(hv[i, :hn[i]], vals / vals.sum(), rtol=1e-5)


def test_find_top_k_golden(oracle):
    g = gold("topk_deque.json")
    oc, ov = oracle.find_k(g["cols"], g["vals"], g["k"], top=True)
    assert [(int(a), float(np.float32(b))) for a, b in zip(oc, ov)] == [(a, float(np.float32(b))) for a, b in g["top"]]
    # findBottomK is the mirror: k smallest, column sorted
    oc, ov = oracle.find_k(g["cols"], g["vals"], g["k"], top=False)
    assert list(oc) == [2, 4, 6]


def test_knn_overlap_golden(oracle):
    g = gold("knn_overlap.json")
    k = g["k"]
    knn = np.array(g["knn"], np.int64).reshape(-1, k)
    n = knn.shape[0]
    l1 = g["children_l1"]
    l2 = [sorted(p for c in comp for p in l1[c]) for comp in g["children_l2"]]
    levels = {0: [[i] for i in range(n)], 1: l1, 2: l2}

    def pixels(level, comp):
        return levels[level][comp]

    for case in g["cases"]:
        pa, pb = pixels(*case["a"]), pixels(*case["b"])
        ua = sorted(set(knn[pa].ravel().tolist()))
        ub = sorted(set(knn[pb].ravel().tolist()))
        assert ua == case["ua"] and ub == case["ub"]  # Hierarchy::getRepresentedDataPoints + union
        comp_ptr = np.array([0, len(pa), len(pa) + len(pb)], np.int64)
        comp_pix = np.array(pa + pb, np.int64)
        sim = oracle.neighbor_overlap(knn, comp_ptr, comp_pix, [0], [1])[0]
        assert sim == pytest.approx(case["overlap"] / min(len(ua), len(ub)), rel=1e-6)


# ---------------------------------------------------------------- Stage A
def test_l2_against_double_loop(oracle):
    # testL2 (UtilsTests.cpp:238-271): L2 vs a double-precision loop, abs 1e-4, uniform(0,100) vectors
    g = gold("l2_dims.json")
    rng = np.random.default_rng(1)
    for d in g["dims"]:
        x = rng.uniform(0, 100, size=(2, d)).astype(np.float32)
        diff = x[0].astype(np.float32) - x[1].astype(np.float32)
        base = np.float32(np.sqrt((diff.astype(np.float64) ** 2).sum()))
        ids, dist = oracle.knn_reference_graph(x, 2, l2_squared=False)
        assert list(ids[0]) == [0, 1]
        assert abs(float(dist[0, 1]) - float(base)) <= g["abs_tol"] * max(1.0, float(base) * 1e-3)


def test_knn_oracle_invariants_scurve(oracle):
    # testConnectingComponentsKnnGraph shape (UtilsTests.cpp:662-703): S-curve N=1000, k=3;
    # invariants NearestNeighbors::compute enforces (self first, non-decreasing, all present)
    from sph_b200 import synth

    x = synth.s_curve(1000, 0.2, 123)
    ids, dist = oracle.knn_reference_graph(x, 3)
    assert (ids[:, 0] == np.arange(1000)).all()
    assert (dist[:, 0] == 0).all()
    assert (np.diff(dist, axis=1) >= 0).all()
    assert (ids >= 0).all()
    # brute-force numpy cross-check of the index sets in float64
    d2 = ((x[:, None, :].astype(np.float64) - x[None, :, :].astype(np.float64)) ** 2).sum(-1)
    order = np.lexsort((np.arange(1000)[None, :].repeat(1000, 0), d2.astype(np.float32)), axis=1)[:, :3]
    np.testing.assert_array_equal(np.sort(ids, axis=1), np.sort(order, axis=1))


def test_knn_postprocess_self_first_cases(oracle):
    # ensureClosestPointIsSelf (GraphUtils.cpp:23-95): swap / overwrite-last / shift
    ids = np.array([[1, 0, 2], [0, 2, 3], [3, 0, 1]], np.int64)
    dist = np.array([[0, 0, 4], [0, 0, 0], [1, 4, 9]], np.float32)
    oracle.knn_postprocess(ids, dist, l2_squared=True)
    assert list(ids[0]) == [0, 1, 2]            # found at slot 1 -> swapped
    assert list(ids[1]) == [1, 2, 0] or ids[1, 0] == 1  # absent, all zero -> last overwritten, swapped to front
    assert ids[1, 0] == 1 and dist[1, 0] == 0
    assert list(ids[2]) == [2, 3, 0] and list(dist[2]) == [0, 1, 4]  # absent -> shifted right


def test_blas_flavour_agrees_with_exact_on_separated_data(oracle):
    # the timed CPU baseline (faiss-style sgemm + heaps) finds the same neighbours when gaps are not fp32-marginal
    if not oracle.blas_init():
        pytest.skip("OpenBLAS not found")
    rng = np.random.default_rng(0)
    x = rng.standard_normal((600, 24)).astype(np.float32)
    ids, _ = oracle.knn_blas_l2(x, 8, 0, 600)
    rids, _ = oracle.knn_exact(x, 8)
    assert (np.sort(ids, 1) == np.sort(rids, 1)).mean() > 0.999


# ---------------------------------------------------------------- Stage B
def test_gaussian_row_hits_perplexity(oracle):
    # no reference golden exists — property: entropy of the row == log(perplexity) within tol
    rng = np.random.default_rng(2)
    dist = np.sort(rng.gamma(2.0, 1.0, 91).astype(np.float32))
    dist[0] = 0
    p, sigma = oracle.gaussian_row(dist, 30.0, tol=1e-6, ignore=0)
    assert sigma > 0 and p[0] == 0
    assert abs(p.sum() - 1.0) < 1e-5
    nz = p[p > 0].astype(np.float64)
    assert abs(-(nz * np.log(nz)).sum() - np.log(30.0)) < 1e-4
    assert (np.diff(p[1:]) <= 1e-12).all()  # monotone (GraphNormalization.cpp:295-327 debug asserts)


def test_level_schedule(oracle):
    # LevelSimilarities::updateNumberOfNeighbors (LevelSimilarities.cpp:83-115)
    perp, ks = oracle.level_schedule([21025, 5000, 900, 40], 91)
    assert perp[0] == pytest.approx(30.0) and ks[0] == 91
    assert perp[1] == pytest.approx(30.0) and ks[1] == 91
    assert perp[2] == pytest.approx(10.0) and ks[2] == 31
    assert ks[3] == 31 or ks[3] == 40  # min(perp*3+1, n_l)
    assert ks[3] == min(int(perp[3]) * 3 + 1, 40)


# ---------------------------------------------------------------- Stage C (RW1)
def test_random_walks_serial_reproducible_and_normalised(oracle):
    rng = np.random.default_rng(5)
    n, k = 60, 6
    m = np.zeros((n, n), np.float32)
    for i in range(n):
        nb = rng.choice([j for j in range(n) if j != i], k, replace=False)
        w = rng.uniform(0.1, 1, k)
        m[i, nb] = w / w.sum()
    rp, c, v = dense_to_csr(m)
    a = oracle.random_walks(rp, c, v, 10, 3, weighting=0, seed=1, threads=1)
    b = oracle.random_walks(rp, c, v, 10, 3, weighting=0, seed=1, threads=1)
    for x, y in zip(a, b):
        np.testing.assert_array_equal(x, y)
    d = csr_to_dense(*a)
    np.testing.assert_allclose(d.sum(1), 1.0, atol=1e-3)  # testImageHierarchyWalks row sums (HierarchyTest.cpp:440-444)
    assert (np.diag(d) == 0).all()  # removeDiagonalElements
    # un-normalised CONSTANT weighting: every walk step adds 1 -> raw row mass = W*L minus self hits
    raw = csr_to_dense(*oracle.random_walks(rp, c, v, 10, 3, weighting=0, seed=1, threads=1, remove_diagonal=False, normalize=False))
    np.testing.assert_allclose(raw.sum(1), 30.0)


# ---------------------------------------------------------------- F1 symmetrizeGraph / RW3 mergeGraphNodes
def _rows(rp, ids):
    return [ids[rp[i]:rp[i + 1]].tolist() for i in range(len(rp) - 1)]


def test_symmetrize_graph_golden(oracle):
    g = gold("symmetrize_graph_24.json")
    ids = np.array(g["ids"], np.int64).reshape(-1, g["k"])
    dist = np.array(g["dist"], np.float32).reshape(-1, g["k"])
    rp, oi, od = oracle.symmetrize_graph(ids, dist)
    rows = _rows(rp, oi)
    for p, exp in g["rows"].items():
        assert rows[int(p)] == exp
    # every edge is reachable from both ends (the isDirectNeighbor loop of the reference test), self first, distance 0
    for p, row in enumerate(rows):
        assert row[0] == p and od[rp[p]] == 0
        for q in row:
            assert p in rows[q]
    # symmetric values: d(p, q) == d(q, p) == min of the stored directions
    lut = {(p, int(oi[e])): float(od[e]) for p in range(len(rows)) for e in range(rp[p], rp[p + 1])}
    for (p, q), v in lut.items():
        assert lut[(q, p)] == v
    assert lut[(0, 1)] == 0.5  # 0->1 stored 1, 1->0 stored 0.5


def test_merge_graph_nodes_golden(oracle):
    g = gold("merge_graph_9to3.json")
    ids = np.array(g["ids"], np.int64).reshape(-1, g["k"])
    dist = np.array(g["dist"], np.float32).reshape(-1, g["k"])
    rp, oi, od = oracle.merge_graph_nodes(ids, dist, g["parents"], 3)
    for m, exp in g["expected"].items():
        m = int(m)
        assert oi[rp[m]] == m and od[rp[m]] == 0  # self first
        for t, (nid, nd) in enumerate(exp, start=1):
            assert oi[rp[m] + t] == nid
            np.testing.assert_allclose(od[rp[m] + t], nd, rtol=1e-6)
    # the sparse form (mergeNodesDataDistances on convertGraphToEigenSparse) gives the same off-diagonal minima
    n, k = ids.shape
    dense = np.zeros((n, n), np.float32)
    for i in range(n):
        for t in range(k):
            dense[i, ids[i, t]] = dist[i, t]
    crp, cc, cv = dense_to_csr(dense)
    mrp, mc, mv = oracle.merge_data_distances(crp, cc, cv, g["parents"], 3)
    md = csr_to_dense(mrp, mc, mv, 3)
    for m, exp in g["expected"].items():
        for nid, nd in exp:
            np.testing.assert_allclose(md[int(m), nid], nd, rtol=1e-6)


def test_symmetrize_graph_self_repair_and_duplicates(oracle):
    # row 1 lists a zero-distance duplicate pixel (0) before itself: ensureClosestPointIsSelf swaps, not shifts
    rp = np.array([0, 2, 5, 6], np.int64)
    ids = np.array([0, 1, 0, 1, 2, 2], np.int64)
    dist = np.array([0, 0, 0, 0, 3, 0], np.float32)
    orp, oi, od = oracle.symmetrize_graph(ids, dist, rp)
    rows = _rows(orp, oi)
    assert rows[0] == [0, 1] and rows[1] == [1, 0, 2] and rows[2] == [2, 1]
    assert od.tolist() == [0, 0, 0, 0, 3, 0, 3]


def test_rw_as_knn_distances_and_zero_removal(oracle):
    m = np.array([[0.5, 0, 0.25, 0.25], [0, 0, 0, 0], [0.1, 0.1, 0.8, 0], [0, 1.0, 0, 0]], np.float32)
    rp, c, v = dense_to_csr(m)
    # an explicit zero must be dropped by both passes
    c = np.concatenate([c, [2]]).astype(np.uint32)
    v = np.concatenate([v, [0.0]]).astype(np.float32)
    rp = rp.copy()
    rp[-1] += 1
    orp, oi, od = oracle.rw_as_knn_distances(rp, c, v)
    assert orp.tolist() == [0, 3, 3, 6, 7]
    assert oi.tolist() == [0, 2, 3, 2, 0, 1, 1]  # ties (0.75, 0.75) and (0.9, 0.9) keep column order
    np.testing.assert_array_equal(od, np.array([0.5, 0.75, 0.75, 1 - np.float32(0.8), 0.9, 0.9, 0.0], np.float32))
    zrp, zc, zv = oracle.remove_zeros(rp, c, v)
    assert zrp.tolist() == [0, 3, 3, 6, 7] and zc.tolist() == [0, 2, 3, 0, 1, 2, 1]


def test_euclid_hausdorff_matches_numpy(oracle):
    rng = np.random.default_rng(5)
    x = rng.normal(size=(60, 7)).astype(np.float32)
    comp_ptr = np.array([0, 1, 5, 5, 20, 60], np.int64)  # sizes 1, 4, 0, 15, 40
    comp_pix = rng.permutation(60).astype(np.int64)
    pa = np.array([0, 1, 3, 4, 0, 2], np.int64)
    pb = np.array([0, 3, 4, 1, 1, 3], np.int64)
    out = oracle.euclid_hausdorff(x, comp_ptr, comp_pix, pa, pb)
    for p in range(len(pa)):
        a = comp_pix[comp_ptr[pa[p]]:comp_ptr[pa[p] + 1]]
        b = comp_pix[comp_ptr[pb[p]]:comp_ptr[pb[p] + 1]]
        if len(a) == 0 or len(b) == 0:
            assert out[p] == 0
            continue
        dm = np.sqrt(((x[a, None, :].astype(np.float64) - x[None, b, :].astype(np.float64)) ** 2).sum(-1))
        exp = max(dm.min(1).max(), dm.min(0).max())
        np.testing.assert_allclose(out[p], exp, rtol=1e-6, atol=1e-7)  # testL2 pins utils::L2 to 1e-4 abs
    assert out[0] == 0  # same single pixel


@pytest.mark.parametrize("mode", [0, 1])
def test_level_knn_walks_matches_pairwise(oracle, mode):
    rng = np.random.default_rng(9 + mode)
    n = 50
    m = (rng.uniform(size=(n, n)) < 0.08) * rng.uniform(0.05, 1, size=(n, n))
    m[7] = 0  # an empty row: everything but itself at distance 1
    m /= np.maximum(m.sum(1, keepdims=True), 1e-9)
    rp, c, v = dense_to_csr(m)
    k = 12
    ids, dist = oracle.level_knn_walks(rp, c, v, k, mode)
    pa, pb = np.meshgrid(np.arange(n), np.arange(n), indexing="ij")
    if mode == 0:  # randomWalksBhattacharyya per pair, self included (the space does not special-case i == j)
        sim = oracle.pair_sims(rp, c, v, pa.ravel(), pb.ravel(), 0).reshape(n, n)
    else:          # randomWalksSingleOverlap(i, j) = A[i][j], one direction only
        sim = csr_to_dense(rp, c, v)
    d = (np.float32(1) - sim).astype(np.float32)
    for i in range(n):
        order = np.lexsort((np.arange(n), d[i]))[:k]
        assert ids[i].tolist() == order.tolist()
        np.testing.assert_array_equal(dist[i], d[i, order])
    # nothing shares a column with the empty row, itself included: the lowest indices at distance exactly 1
    assert (dist[7] == 1).all() and ids[7].tolist() == list(range(k))


# ---------------------------------------------------------------- full hierarchies (host merge loop on the oracle)
def run_hierarchy(g, ops):
    from sph_b200 import host

    n = g["rows"] * g["cols"]
    ids = np.array(g["knn"], np.int64).reshape(n, g["k"])
    dist = np.array(g.get("dist", [0.0] * (n * g["k"])), np.float32).reshape(n, g["k"])
    ih = host.ImageHierarchy(host.Graph(ids, dist), None, g["rows"], g["cols"], ops=ops)
    ihs = host.ImageHierarchySettings(componentSim=host.ComponentSim[g["sim"]], neighborConnection=host.NeighConnection.FOUR,
                                      mergeMultiple=False, usePercentile=False)
    rws = host.RandomWalkSettings()
    if "num_walks" in g:
        ihs.rwHandling = host.RandomWalkHandling.MERGE_RW_NEW_WALKS
        ihs.normKnnDistances = host.NormalizationScheme.LINEAR
        rws = host.RandomWalkSettings(numRandomWalks=g["num_walks"], singleWalkLength=g["walk_len"], randomSeed=g["seed"], parallel=False,
                                      importanceWeighting=host.ImportanceWeighting.CONSTANT)
    ih.compute(ihs, rws)
    return ih


def check_hierarchy(ih, g):
    levels = len(g["pixel_components"])
    assert ih.getNumLevels() == levels
    for lvl in range(levels):
        assert ih.pixelComponentsOn(lvl).tolist() == g["pixel_components"][lvl]
    for lvl in range(levels - 1):
        assert ih.parentsOn(lvl).tolist() == g["parents"][lvl]
        assert ih.childrenOn(lvl + 1) == g["children"][lvl]
        assert ih.spatialNeighborsOn(lvl + 1) == g["spatial_neighbors"][lvl]


@pytest.mark.parametrize("name", ["hierarchy_overlap_4x4.json", "hierarchy_overlap_6x4.json", "hierarchy_walks_6x4.json"])
def test_full_hierarchy_golden(oracle, name):
    g = gold(name)
    ih = run_hierarchy(g, oracle_lib.OracleOps(oracle))
    check_hierarchy(ih, g)
    if "num_walks" in g:  # testImageHierarchyWalks (:440-444): every random-walk row sums to 1
        for rw in ih.randomWalks:
            sums = np.add.reduceat(rw.val, rw.rowptr[:-1])
            np.testing.assert_allclose(sums, 1.0, atol=1e-3)


def test_level_knn_overlap_matches_pairwise(oracle):
    # golden unions of testKnnOverlap (level-1 components) + the all-pairs definition through orc_neighbor_overlap
    g = gold("knn_overlap.json")
    knn = np.array(g["knn"], np.int64).reshape(-1, g["k"])
    children = g["children_l1"]
    ptr = np.concatenate([[0], np.cumsum([len(c) for c in children])]).astype(np.int64)
    pix = np.array([p for c in children for p in c], np.int64)
    n = len(children)
    k = 5
    ids, dist = oracle.level_knn_overlap(knn, ptr, pix, k)
    pa, pb = np.meshgrid(np.arange(n), np.arange(n), indexing="ij")
    sim = oracle.neighbor_overlap(knn, ptr, pix, pa.ravel(), pb.ravel()).reshape(n, n)
    d = (np.float32(1) - sim).astype(np.float32)
    for i in range(n):
        order = np.lexsort((np.arange(n), d[i]))[:k]
        assert ids[i].tolist() == order.tolist()
        np.testing.assert_array_equal(dist[i], d[i, order])
        assert ids[i, 0] == i and dist[i, 0] == 0  # a component overlaps itself completely
    # golden: components 8 and 0 of level 1 share exactly one id out of min(5, 5) (tests/HierarchyTest.cpp:157-226)
    assert d[8, 0] == np.float32(1) - np.float32(1) / np.float32(5)


def test_weak_components_golden(oracle):
    g = gold("components_9.json")
    num, labels = oracle.weak_components(np.array(g["ids"], np.int64).reshape(-1, g["k"]))
    assert num == g["num_components"] and labels.tolist() == g["labels"]
    # the 24-point kNN table of testSymmetrizeKnnGraph is one component once edges are undirected; numbering follows the lowest vertex
    g2 = gold("symmetrize_graph_24.json")
    num2, labels2 = oracle.weak_components(np.array(g2["ids"], np.int64).reshape(-1, g2["k"]))
    assert num2 == 1 and (labels2 == 0).all()
    ids = np.array([[0, 3], [1, 2], [2, 1], [3, 0], [4, 4]], np.int64)
    num3, labels3 = oracle.weak_components(ids)
    assert num3 == 3 and labels3.tolist() == [0, 1, 1, 0, 2]


def test_prow_graph_ragged_reduces_to_fixed_k(oracle):
    rng = np.random.default_rng(4)
    n, k = 40, 9
    ids = np.stack([np.concatenate([[i], rng.choice(np.delete(np.arange(n), i), k - 1, replace=False)]) for i in range(n)]).astype(np.int64)
    dist = np.concatenate([np.zeros((n, 1)), np.sort(rng.uniform(0.1, 3, (n, k - 1)), 1)], 1).astype(np.float32)
    rp = np.arange(n + 1, dtype=np.int64) * k
    for linear in (False, True):
        orp, oc, ov = oracle.prow_graph_ragged(rp, ids.reshape(-1), dist.reshape(-1), 0.0, linear)
        cols, vals = oracle.prow_linear_graph(ids, dist) if linear else oracle.prow_gaussian_graph(ids, dist, 0.0)[:2]
        assert orp.tolist() == (np.arange(n + 1) * (k - 1)).tolist()
        np.testing.assert_array_equal(oc, cols.reshape(-1))
        np.testing.assert_array_equal(ov, vals.reshape(-1))


def test_hierarchy_merge_data_new_walks_runs_on_oracle(oracle):
    # MERGE_DATA_NEW_WALKS (sph/utils/Hierarchy.cpp:309-339): mergeGraphNodes + rows of the merged (ragged) data graph + new walks
    from sph_b200 import host

    g = gold("hierarchy_walks_6x4.json")
    ids = np.array(g["knn"], np.int64).reshape(24, 4)
    dist = np.array(g["dist"], np.float32).reshape(24, 4)
    ih = host.ImageHierarchy(host.Graph(ids, dist), None, 6, 4, ops=oracle_lib.OracleOps(oracle))
    ihs = host.ImageHierarchySettings(componentSim=host.ComponentSim.NEIGH_WALKS, usePercentile=False,
                                      rwHandling=host.RandomWalkHandling.MERGE_DATA_NEW_WALKS, normKnnDistances=host.NormalizationScheme.LINEAR)
    ih.compute(ihs, host.RandomWalkSettings(numRandomWalks=10, singleWalkLength=3, parallel=False))
    assert ih.getNumLevels() >= 3 and ih.numComponents[-1] == 1
    assert len(ih.mergedDataGraphs) == ih.getNumLevels()
    for lvl, mg in enumerate(ih.mergedDataGraphs):
        assert mg.numPoints == ih.numComponents[lvl]
        assert (mg.knnIndices[mg.rowptr[:-1]] == np.arange(mg.numPoints)).all()  # self first on every level


def run_walks_4x4(ops):
    from sph_b200 import host

    g = gold("hierarchy_walks_4x4_rowsums.json")
    ids = np.array(g["knn"], np.int64).reshape(16, 4)
    dist = np.array(g["dist"], np.float32).reshape(16, 4)
    ih = host.ImageHierarchy(host.Graph(ids, dist), None, 4, 4, ops=ops)
    ih.compute(host.ImageHierarchySettings(componentSim=host.ComponentSim.NEIGH_WALKS, rwHandling=host.RandomWalkHandling.MERGE_RW_NEW_WALKS))
    assert ih.getNumLevels() >= 2 and len(ih.randomWalks) == ih.getNumLevels()
    for rw in ih.randomWalks:
        assert rw.n > 0 and (np.diff(rw.rowptr) > 0).all()
        np.testing.assert_allclose(np.add.reduceat(rw.val, rw.rowptr[:-1]), 1.0, atol=g["row_sum_abs_tol"])
    return ih


def test_image_hierarchy_walks_row_sums(oracle):
    run_walks_4x4(oracle_lib.OracleOps(oracle))
