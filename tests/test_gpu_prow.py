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
