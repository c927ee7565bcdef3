"""GPU parity tests, Stage B: perplexity rows (through the C ABI) vs the CPU oracle.

Tolerance: 1e-5 relative on probabilities (north_star); columns bit-exact.  The GPU
runs the same fp64 bisection; only the order of the fp64 row sums differs.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

RTOL = 1e-5


def _graph(n, k, seed, scale=1.0, zero_rows=0):
    rng = np.random.default_rng(seed)
    dist = np.sort(rng.gamma(2.0, scale, size=(n, k)).astype(np.float32), axis=1)
    dist[:, 0] = 0.0
    ids = np.empty((n, k), np.int64)
    for i in range(n):
        others = rng.choice(max(n, k + 1) * 3, size=k - 1, replace=False)
        ids[i, 0] = i
        ids[i, 1:] = others
    if zero_rows:
        dist[:zero_rows] = 0.0
    return ids, dist


@pytest.mark.parametrize("k,perp", [(91, 0.0), (91, 30.0), (31, 10.0), (301, 100.0), (16, 0.0), (130, 40.0)])
def test_gaussian_graph(capi, oracle, k, perp):
    ids, dist = _graph(400, k, seed=k)
    cols, vals, sigma = capi.prow_gaussian_graph(ids, dist, perp, want_sigma=True)
    rcols, rvals, rsigma = oracle.prow_gaussian_graph(ids, dist, perp)
    np.testing.assert_array_equal(cols, rcols)
    np.testing.assert_allclose(vals, rvals, rtol=RTOL, atol=1e-12)
    np.testing.assert_allclose(sigma, rsigma, rtol=1e-9)
    np.testing.assert_allclose(vals.sum(axis=1), 1.0, atol=1e-3)


def test_gaussian_graph_fallbacks(capi, oracle):
    # all-zero rows -> uniform; huge distances -> sigma < 1e-3 -> inverted-linear fallback;
    # tiny spread -> search fails -> uniform then fallback
    ids, dist = _graph(64, 31, seed=1, zero_rows=8)
    dist[8:16] *= 1e7
    dist[16:24, 1:] = 5.0
    dist[24:32] *= 1e-9
    cols, vals, sigma = capi.prow_gaussian_graph(ids, dist, 10.0, want_sigma=True)
    rcols, rvals, rsigma = oracle.prow_gaussian_graph(ids, dist, 10.0)
    np.testing.assert_array_equal(cols, rcols)
    np.testing.assert_allclose(vals, rvals, rtol=RTOL, atol=1e-12)
    np.testing.assert_allclose(sigma, rsigma, rtol=1e-9)


def test_linear_graph(capi, oracle):
    ids, dist = _graph(300, 25, seed=3)
    cols, vals = capi.prow_linear_graph(ids, dist)
    rcols, rvals = oracle.prow_linear_graph(ids, dist)
    np.testing.assert_array_equal(cols, rcols)
    np.testing.assert_allclose(vals, rvals, rtol=1e-6)


def test_gaussian_inplace_ragged(capi, oracle):
    rng = np.random.default_rng(11)
    lens = rng.integers(0, 200, size=300)
    lens[:5] = [0, 1, 2, 3, 301]
    rowptr = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    val = rng.gamma(2.0, 1.0, size=int(rowptr[-1])).astype(np.float32)
    # a row whose tail underflows below 1e-10 exercises the write-back cursor quirk
    b, e = rowptr[10], rowptr[11]
    if e - b > 4:
        val[b:e] = np.linspace(0.0, 400.0, e - b, dtype=np.float32)
    for perp, ignore in [(0.0, -1), (12.0, -1), (0.0, 0)]:
        out = capi.prow_gaussian_inplace(rowptr, val, perp, ignore)
        ref = oracle.prow_gaussian_inplace(rowptr, val, perp, ignore)
        np.testing.assert_allclose(out, ref, rtol=RTOL, atol=1e-12)


def test_gaussian_flat(capi, oracle):
    _, dist = _graph(200, 91, seed=5)
    dist = dist ** 2
    out = capi.prow_gaussian_flat(dist, 30.0)
    ref = oracle.prow_gaussian_flat(dist, 30.0)
    np.testing.assert_allclose(out, ref, rtol=RTOL, atol=1e-12)


def test_search_ends_at_the_reference_beta(capi, oracle):
    """The search screens every iterate in fp32 (with a proven error band) before falling back to the reference's fp64 evaluation:
    every branch it takes must be the reference's, so the final beta — returned as sigma = sqrt(1 / 2 beta) — is bit-identical.
    20 000 rows over 12 decades of distance scale, large offsets (beta * d_min beyond the fp32 underflow of p), near-constant rows,
    heavy ties, perplexities from 2 to k - 2."""
    rng = np.random.default_rng(123)
    n, k = 20000, 91
    base = np.sort(rng.gamma(2.0, 1.0, size=(n, k)).astype(np.float32), axis=1)
    scale = (10.0 ** rng.uniform(-6, 6, size=(n, 1))).astype(np.float32)
    dist = base * scale
    off = np.where(rng.uniform(size=(n, 1)) < 0.3, 10.0 ** rng.uniform(-3, 3, size=(n, 1)), 0.0).astype(np.float32)
    dist = dist + off * scale
    dist[:2000] = np.round(dist[:2000] / scale[:2000] * 2) / 2 * scale[:2000]      # ties
    dist[2000:3000, 1:] = dist[2000:3000, 1:2] * (1 + 1e-6 * rng.uniform(size=(1000, k - 1))).astype(np.float32)  # near-constant
    dist = np.sort(dist, axis=1)
    dist[:, 0] = 0
    offs = 1 + np.argsort(rng.random((n, 400)), axis=1)[:, :k - 1]   # distinct neighbours per row, as in a kNN graph
    ids = np.concatenate([np.arange(n)[:, None], (np.arange(n)[:, None] + offs) % n], axis=1).astype(np.int64)
    for perp in (0.0, 2.0, 50.0, 89.0):
        cols, vals, sigma = capi.prow_gaussian_graph(ids, dist, perp, want_sigma=True)
        rcols, rvals, rsigma = oracle.prow_gaussian_graph(ids, dist, perp)
        np.testing.assert_array_equal(sigma, rsigma)
        np.testing.assert_array_equal(cols, rcols)
        np.testing.assert_allclose(vals, rvals, rtol=RTOL, atol=1e-12)


# ---- UMAP-flavoured rows (utils::computeExponentialDistributions, sph/utils/GraphNormalization.cpp:413-593) ----
# fp32 search on both sides (umappp's Float_ is float); the device sums in warp-tree order and uses CUDA's expf, so a search may stop
# one bisection step apart from the oracle's: |sum - target| < 1e-5 on a sum of log2(k) bounds the effect on a weight well below 1e-4.
URTOL = 1e-4


@pytest.mark.parametrize("k,sort_cols", [(91, False), (91, True), (16, False), (31, True), (130, False), (301, True), (1025, False)])
def test_exponential_graph(capi, oracle, k, sort_cols):
    ids, dist = _graph(300, k, seed=1000 + k, scale=0.5 if k % 2 else 3.0)
    cols, vals = capi.prow_exponential_graph(ids, dist, sort_cols)
    rcols, rvals = oracle.prow_exponential_graph(ids, dist, sort_cols)
    np.testing.assert_array_equal(cols, rcols)
    np.testing.assert_allclose(vals, rvals, rtol=URTOL, atol=1e-7)
    np.testing.assert_allclose(vals.sum(axis=1), np.log2(k), atol=3e-5 * k)  # the defining equation of the rows


def test_exponential_graph_edge_rows(capi, oracle):
    k = 9
    ids = np.tile(np.arange(k, dtype=np.int64), (6, 1))
    dist = np.zeros((6, k), np.float32)
    dist[1, 1:] = 2.5
    dist[2, 1:] = [0, 1, 2, 3, 4, 5, 6, 7]
    dist[3, 1:] = [1, 1, 1, 1, 1, 1, 1, 1e6]
    dist[4, 1:] = [1e-3, 1, 1, 1, 1, 1, 1, 1]
    dist[5, 1:] = [0, 0, 0, 0, 0, 0, 0, 4]
    cols, vals = capi.prow_exponential_graph(ids, dist)
    rcols, rvals = oracle.prow_exponential_graph(ids, dist)
    np.testing.assert_array_equal(cols, rcols)
    np.testing.assert_allclose(vals, rvals, rtol=URTOL, atol=1e-7)
    assert (vals[0] == 1).all() and (vals[1] == 1).all() and (vals[5] == 1).all() and vals[3, 7] == 0.0


def test_exponential_inplace(capi, oracle):
    rng = np.random.default_rng(77)
    lens = np.concatenate([rng.integers(0, 130, size=300), [0, 1, 2, 128, 129, 320, 321, 1024]])
    rowptr = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    col = np.concatenate([np.sort(rng.choice(4096, size=m, replace=False)) for m in lens]).astype(np.uint32)
    val = rng.gamma(2.0, 1.0, size=rowptr[-1]).astype(np.float32)
    val[rng.random(val.size) < 0.02] = 0.0  # stored zeros in the middle of a row
    got = capi.prow_exponential_inplace(rowptr, val)
    ref = oracle.prow_exponential_inplace(rowptr, val)
    np.testing.assert_allclose(got, ref, rtol=URTOL, atol=1e-7)
    got7 = capi.prow_exponential_inplace(rowptr, val, col=col, ignore=7)
    ref7 = oracle.prow_exponential_inplace(rowptr, val, col=col, ignore=7)
    np.testing.assert_allclose(got7, ref7, rtol=URTOL, atol=1e-7)
    assert (got7[col == 7] == val[col == 7]).all() and (col == 7).sum() > 0


def test_exponential_rows_feed_symmetrize_umap(capi, oracle):
    """The UMAP branch of LevelSimilarities::computeProbDist (sph/LevelSimilarities.cpp:499-503 + symmetrizeOutput :566-600):
    exponential rows, then the fuzzy union a + b - ab."""
    n, k = 500, 16
    rng = np.random.default_rng(3)
    x = rng.normal(size=(n, 6)).astype(np.float32)
    ids, dist = capi.knn_exact(x, k)
    cols, vals = capi.prow_exponential_graph(ids, dist, sort_cols=True)
    rowptr = np.arange(n + 1, dtype=np.int64) * (k - 1)
    got = capi.symmetrize(rowptr, cols.ravel(), vals.ravel(), umap=True)
    ref = oracle.symmetrize(rowptr, cols.ravel(), vals.ravel(), umap=True)
    for g, r in zip(got, ref):
        np.testing.assert_array_equal(g, r)
