"""GPU parity tests, Stages B6 / C / D: CSR graph passes (through the C ABI) vs the CPU oracle and the
reference's golden vectors (tests/golden)."""
import json
from pathlib import Path

import numpy as np
import pytest

from oracle_lib import csr_to_dense, dense_to_csr

pytestmark = pytest.mark.gpu
GOLD = Path(__file__).resolve().parent / "golden"


def gold(name):
    return json.loads((GOLD / name).read_text())


def random_stochastic(n, nnz_per_row, seed, diag=False):
    rng = np.random.default_rng(seed)
    rowptr = [0]
    col, val = [], []
    for i in range(n):
        m = int(rng.integers(1, nnz_per_row + 1))
        pool = np.arange(n) if diag else np.delete(np.arange(n), i)
        c = np.sort(rng.choice(pool, size=min(m, len(pool)), replace=False))
        w = rng.uniform(0.05, 1.0, len(c))
        col.extend(c.tolist())
        val.extend((w / w.sum()).tolist())
        rowptr.append(len(col))
    return np.array(rowptr, np.int64), np.array(col, np.uint32), np.array(val, np.float32)


# ---------------------------------------------------------------- B6
@pytest.mark.parametrize("umap", [False, True])
def test_symmetrize_golden(capi, umap):
    g = gold("symmetrize_5x5.json")
    rp, c, v = dense_to_csr(g["input"])
    orp, oc, ov = capi.symmetrize(rp, c, v, umap=umap)
    out = csr_to_dense(orp, oc, ov)
    np.testing.assert_allclose(out, np.array(g["umap" if umap else "tsne"], np.float32), atol=g["tol"])


@pytest.mark.parametrize("umap", [False, True])
def test_symmetrize_matches_oracle(capi, oracle, umap):
    rp, c, v = random_stochastic(700, 30, seed=4, diag=True)
    # explicit zeros and an empty row are part of the pattern semantics
    v[::17] = 0.0
    a = capi.symmetrize(rp, c, v, umap=umap)
    b = oracle.symmetrize(rp, c, v, umap=umap)
    np.testing.assert_array_equal(a[0], b[0])
    np.testing.assert_array_equal(a[1], b[1])
    np.testing.assert_array_equal(a[2], b[2])  # one fp32 operation per entry: bit-exact


def test_symmetrize_level0_prow_shape(capi, oracle):
    # the P rows of a kNN graph (fixed k-1 entries per row), as symmetrizeOutput sees them (LevelSimilarities.cpp:589-623)
    rng = np.random.default_rng(9)
    n, k = 3000, 31
    ids = np.stack([np.concatenate([[i], rng.choice(np.delete(np.arange(n), i), k - 1, replace=False)]) for i in range(n)]).astype(np.int64)
    dist = np.sort(rng.gamma(2.0, 1.0, size=(n, k)).astype(np.float32), axis=1)
    dist[:, 0] = 0
    cols, vals = capi.prow_gaussian_graph(ids, dist, 10.0)
    rp = np.arange(0, n * (k - 1) + 1, k - 1, dtype=np.int64)
    a = capi.symmetrize(rp, cols.ravel(), vals.ravel())
    b = oracle.symmetrize(rp, cols.ravel(), vals.ravel())
    for x, y in zip(a, b):
        np.testing.assert_array_equal(x, y)
    d = csr_to_dense(*a)
    np.testing.assert_array_equal(d, d.T)


# ---------------------------------------------------------------- RW2
def test_merge_nodes_synth_golden(capi):
    g = gold("merge_nodes_synth.json")
    rp, c, v = dense_to_csr(g["input"])
    orp, oc, ov = capi.rw_merge(rp, c, v, g["parents"], 3, norm=False, weight_by_size=False)
    np.testing.assert_array_equal(csr_to_dense(orp, oc, ov), np.array(g["expected"], np.float32))


@pytest.mark.parametrize("norm,wbs", [(True, True), (False, True), (True, False), (False, False)])
def test_merge_matches_oracle(capi, oracle, norm, wbs):
    n, nm = 2000, 333
    rp, c, v = random_stochastic(n, 40, seed=7)
    rng = np.random.default_rng(8)
    parents = rng.integers(0, nm, size=n)
    parents[:nm] = np.arange(nm)
    a = capi.rw_merge(rp, c, v, parents, nm, norm=norm, weight_by_size=wbs)
    b = oracle.rw_merge(rp, c, v, parents, nm, norm=norm, weight_by_size=wbs)
    np.testing.assert_array_equal(a[0], b[0])
    np.testing.assert_array_equal(a[1], b[1])
    np.testing.assert_array_equal(a[2], b[2])  # same fp32 accumulation order as the single-thread reference


# ---------------------------------------------------------------- D4
def test_pair_sims_match_oracle(capi, oracle):
    rp, c, v = random_stochastic(1500, 60, seed=12)
    rng = np.random.default_rng(13)
    pa = rng.integers(0, 1500, size=5000)
    pb = rng.integers(0, 1500, size=5000)
    pb[:10] = pa[:10]
    for mode in (0, 1):
        np.testing.assert_array_equal(capi.pair_sims(rp, c, v, pa, pb, mode), oracle.pair_sims(rp, c, v, pa, pb, mode))


# ---------------------------------------------------------------- D1
def _check_sims(capi, oracle, rp, c, v, k, prune, variant, comp_sizes=None):
    gc, gv, gn = capi.create_similarities(rp, c, v, k, prune, variant, comp_sizes)
    oc, ov, on = oracle.create_similarities(rp, c, v, k, prune, variant, comp_sizes)
    n = len(rp) - 1
    bad = 0
    for i in range(n):
        if gn[i] != on[i] or not np.array_equal(gc[i, :gn[i]], oc[i, :on[i]]):
            # a different float summation order (L2 atomics vs column-ordered merge dot) may swap two candidates
            # whose distances agree to ~1 ulp at the k-th slot; anything else is a failure
            assert gn[i] == on[i], f"row {i}: nnz {gn[i]} vs {on[i]}"
            diff = set(gc[i, :gn[i]].tolist()) ^ set(oc[i, :on[i]].tolist())
            assert len(diff) <= 2, f"row {i}: selected sets differ by {diff}"
            bad += 1
            continue
        np.testing.assert_allclose(gv[i, :gn[i]], ov[i, :on[i]], rtol=2e-5, atol=2e-6)
        assert (gc[i, gn[i]:] == 0xFFFFFFFF).all()
    assert bad <= max(1, n // 200)


def test_similarities_golden_5x5(capi, oracle):
    g = gold("sims_5x5.json")
    rp, c, v = dense_to_csr(g["input"])
    for variant in (1, 2):
        _check_sims(capi, oracle, rp, c, v, g["k"], g["prune"], variant)


@pytest.mark.parametrize("variant,k", [(1, 31), (2, 31), (1, 91), (0, 10)])
def test_similarities_match_oracle(capi, oracle, variant, k):
    # rows shaped like merged random walks: a few hundred columns drawn from a local window
    rng = np.random.default_rng(k + variant)
    n = 1200
    rowptr = [0]
    col, val = [], []
    for i in range(n):
        m = int(rng.integers(5, 120))
        c = np.unique(np.clip(i + rng.integers(-150, 150, size=m), 0, n - 1))
        w = rng.gamma(0.7, 1.0, len(c)) + 1e-6
        col.extend(c.tolist())
        val.extend((w / w.sum()).tolist())
        rowptr.append(len(col))
    rp, c, v = np.array(rowptr, np.int64), np.array(col, np.uint32), np.array(val, np.float32)
    _check_sims(capi, oracle, rp, c, v, k, 1e-4, variant)
    sizes = rng.integers(1, 50, size=n)
    _check_sims(capi, oracle, rp, c, v, k, 1e-4, variant, comp_sizes=sizes)


# ---------------------------------------------------------------- RW1
def _ring(n, deg):
    """Every node steps to one of its `deg` successors on a ring with equal probability."""
    rowptr = np.arange(0, n * deg + 1, deg, dtype=np.int64)
    col = np.sort(((np.arange(n)[:, None] + np.arange(1, deg + 1)[None, :]) % n).astype(np.uint32), axis=1)
    val = np.full(n * deg, 1.0 / deg, np.float32)
    return rowptr, col.ravel(), val


def test_random_walks_deterministic_chain_is_exact(capi, oracle):
    # one successor per node: the walk is deterministic, so the CUDA result must equal the reference bit for bit
    n = 500
    rp, c, v = _ring(n, 1)
    for weighting in (capi.W_CONSTANT, capi.W_LINEAR, capi.W_NORMAL, capi.W_ONLYLAST, capi.W_FIRST_VISIT):
        a = capi.random_walks(rp, c, v, 7, 5, weighting, seed=3)
        b = oracle.random_walks(rp, c, v, 7, 5, weighting, seed=3)
        np.testing.assert_array_equal(a[0], b[0])
        np.testing.assert_array_equal(a[1], b[1])
        np.testing.assert_allclose(a[2], b[2], rtol=1e-6)


def test_random_walks_invariants(capi):
    rp, c, v = random_stochastic(800, 12, seed=21)
    W, L = 90, 15
    orp, oc, ov = capi.random_walks(rp, c, v, W, L, capi.W_CONSTANT, seed=5, remove_diagonal=False, normalize=False)
    d = csr_to_dense(orp, oc, ov)
    np.testing.assert_array_equal(d.sum(1), np.full(800, W * L, np.float32))  # every step adds weight 1
    for i in range(800):
        assert (np.diff(oc[orp[i]:orp[i + 1]].astype(np.int64)) > 0).all()
    orp, oc, ov = capi.random_walks(rp, c, v, W, L, capi.W_NORMAL, seed=5)
    d = csr_to_dense(orp, oc, ov)
    np.testing.assert_allclose(d.sum(1), 1.0, atol=1e-3)   # HierarchyTest.cpp:440-444
    assert (np.diag(d) == 0).all()
    # same seed -> same result, independent of launch geometry; different seed -> different walks
    again = capi.random_walks(rp, c, v, W, L, capi.W_NORMAL, seed=5)
    np.testing.assert_array_equal(again[2], ov)
    other = capi.random_walks(rp, c, v, W, L, capi.W_NORMAL, seed=6)
    assert not np.array_equal(other[2], ov) or not np.array_equal(other[1], oc)


@pytest.mark.parametrize("weighting", [0, 1, 2])
def test_random_walks_statistical_parity(capi, oracle, weighting):
    """Ring with 4 equiprobable successors: the expected (un-normalised) hit mass at offset o is the same for every
    start point, E[o] = W * sum_step w(step) * P(S_step = o).  Averaged over all start points both the CUDA sampler
    (Philox) and the reference sampler (mt19937_64) must agree with the exact expectation within 6 sigma, and
    with each other."""
    n, deg, W, L = 2000, 4, 30, 6
    rp, c, v = _ring(n, deg)
    # exact distribution of the offset after each step
    step_dist = np.zeros((L, L * deg + 1))
    cur = np.zeros(L * deg + 1)
    cur[0] = 1.0
    for s in range(L):
        nxt = np.zeros_like(cur)
        for j in range(1, deg + 1):
            nxt[j:] += cur[:-j] / deg
        cur = nxt
        step_dist[s] = cur
    if weighting == 0:
        w = np.ones(L)
    elif weighting == 1:
        w = np.array([np.float32(1.0 - s / L) for s in range(L)], np.float64)
    else:
        w = np.array([np.exp(np.float32(-0.5) * np.float32(s * 3.0 / L) ** 2) for s in range(L)], np.float64)
    expect = W * (w[:, None] * step_dist).sum(0)
    var = W * ((w[:, None] ** 2) * step_dist * (1 - step_dist)).sum(0)  # upper bound (steps of one walk are positively dependent at most)

    def offsets(res):
        orp, oc, ov = res
        acc = np.zeros(L * deg + 1)
        rows = np.repeat(np.arange(n), np.diff(orp))
        off = (oc.astype(np.int64) - rows) % n
        np.add.at(acc, off, ov.astype(np.float64))
        return acc / n

    gpu = offsets(capi.random_walks(rp, c, v, W, L, weighting, seed=11, remove_diagonal=False, normalize=False))
    ref = offsets(oracle.random_walks(rp, c, v, W, L, weighting, seed=11, threads=1, remove_diagonal=False, normalize=False))
    sigma = np.sqrt(np.maximum(var, 1e-12) * L / n)  # generous: whole-walk dependence
    assert np.all(np.abs(gpu - expect) <= 6 * sigma + 1e-9), np.abs(gpu - expect) / sigma
    assert np.all(np.abs(ref - expect) <= 6 * sigma + 1e-9)
    assert abs(gpu.sum() - expect.sum()) < 1e-6 * expect.sum()


# ---------------------------------------------------------------- D3
def test_find_k_golden_and_oracle(capi, oracle):
    g = gold("topk_deque.json")
    rp = np.array([0, len(g["cols"])], np.int64)
    oc, ov, nnz = capi.find_k(rp, g["cols"], g["vals"], g["k"], top=True)
    assert [(int(a), float(b)) for a, b in zip(oc[0, :nnz[0]], ov[0, :nnz[0]])] == [(a, float(np.float32(b))) for a, b in g["top"]]
    # ragged random rows incl. ties, zeros and an empty row, both directions
    rng = np.random.default_rng(3)
    lens = rng.integers(0, 200, size=400)
    lens[0] = 0
    rowptr = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    col = np.concatenate([np.sort(rng.choice(5000, size=m, replace=False)) for m in lens]).astype(np.uint32)
    val = (rng.integers(0, 12, size=int(rowptr[-1])) / 11.0).astype(np.float32)  # heavy ties, exact zeros
    for top in (True, False):
        for k in (5, 31):
            oc, ov, nnz = capi.find_k(rowptr, col, val, k, top)
            for i in range(len(lens)):
                rc, rv = oracle.find_k(col[rowptr[i]:rowptr[i + 1]], val[rowptr[i]:rowptr[i + 1]], k, top)
                assert nnz[i] == len(rc)
                np.testing.assert_array_equal(oc[i, :nnz[i]], rc)
                np.testing.assert_array_equal(ov[i, :nnz[i]], rv)


# ---------------------------------------------------------------- D5
def test_neighbor_overlap_golden(capi):
    g = gold("knn_overlap.json")
    k = g["k"]
    knn = np.array(g["knn"], np.int64).reshape(-1, k)
    n = knn.shape[0]
    l1 = g["children_l1"]
    l2 = [sorted(p for c in comp for p in l1[c]) for comp in g["children_l2"]]
    levels = {0: [[i] for i in range(n)], 1: l1, 2: l2}
    for case in g["cases"]:
        pa, pb = levels[case["a"][0]][case["a"][1]], levels[case["b"][0]][case["b"][1]]
        comp_ptr = np.array([0, len(pa), len(pa) + len(pb)], np.int64)
        sim = capi.neighbor_overlap(knn, comp_ptr, np.array(pa + pb, np.int64), [0], [1])[0]
        assert sim == np.float32(np.float32(case["overlap"]) / np.float32(min(len(case["ua"]), len(case["ub"]))))


def test_neighbor_overlap_matches_oracle(capi, oracle):
    rng = np.random.default_rng(6)
    n, k, ncomp = 6000, 31, 500
    knn = np.stack([np.concatenate([[i], np.clip(i + rng.integers(-400, 400, size=k - 1), 0, n - 1)]) for i in range(n)]).astype(np.int64)
    label = np.sort(rng.integers(0, ncomp, size=n))
    label[:ncomp] = np.arange(ncomp)
    order = np.argsort(label, kind="stable")
    comp_ptr = np.concatenate([[0], np.cumsum(np.bincount(label, minlength=ncomp))]).astype(np.int64)
    pa = rng.integers(0, ncomp, size=3000)
    pb = np.clip(pa + rng.integers(-3, 4, size=3000), 0, ncomp - 1)
    a = capi.neighbor_overlap(knn, comp_ptr, order, pa, pb)
    b = oracle.neighbor_overlap(knn, comp_ptr, order, pa, pb)
    np.testing.assert_array_equal(a, b)


# ---------------------------------------------------------------- RW1: binary-search kernel == serial-scan kernel
@pytest.mark.parametrize("weighting", [0, 1, 2, 3])
def test_random_walks_search_equals_scan(capi, weighting):
    """The step of the production kernel is a binary search in the rows' fp64 running sums; the serial scan of the reference
    (`inc += val[e]; if (u < inc)`, SparseMatrixAlgorithms.cpp:129-135) picks the same entry for the same uniform, so the two
    kernels must agree bit for bit — rows of 1..60 entries, an empty row (walk stays put) and rows that do not sum to 1."""
    rp, c, v = random_stochastic(1500, 60, seed=33)
    v = v.copy()
    v[rp[7]:rp[8]] *= 0.5                      # defective row: u beyond the last running sum -> stays put
    lens = np.diff(rp)
    keep = np.ones(len(c), bool)
    keep[rp[11]:rp[12]] = False                # empty row
    rp2 = np.concatenate([[0], np.cumsum(np.where(np.arange(1500) == 11, 0, lens))]).astype(np.int64)
    c2, v2 = c[keep], v[keep]
    W, L = 90, 15
    a = capi.random_walks(rp2, c2, v2, W, L, weighting, seed=3)
    capi.force_walk_scan(True)
    try:
        b = capi.random_walks(rp2, c2, v2, W, L, weighting, seed=3)
    finally:
        capi.force_walk_scan(False)
    for x, y in zip(a, b):
        np.testing.assert_array_equal(x, y)
