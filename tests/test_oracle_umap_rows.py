"""CPU tests of the oracle's UMAP-flavoured rows (utils::computeExponentialDistributions,
sph/utils/GraphNormalization.cpp:413-593).  umappp 3.2.0 (the library the reference calls) is not part of the reference tree and
the reference holds no golden vector for these rows, so parity is UNPINNED: what is checked here is the published definition
itself — rho, the defining equation sum_j w_j = log2(n_neighbours + 1), the min-sigma clamp and the edge cases — against an
independent numpy statement written from the same definition in fp64 with a root finder of its own.
"""
import numpy as np
import pytest


def _numpy_row(d):
    """fp64 restatement with a plain bisection on [1e-30, 1e30] to 1e-14: the ideal answer the fp32 search approaches."""
    d = np.asarray(d, np.float64)
    nz = np.nonzero(d != 0)[0]
    rho = d[nz[0]] if nz.size else 0.0
    act = d > rho
    if not act.any():
        return np.ones_like(d)
    target = np.log2(len(d) + 1.0)
    f = lambda s: (~act).sum() + np.exp(-(d[act] - rho) / s).sum() - target
    lo, hi = 1e-30, 1e30
    for _ in range(400):
        mid = np.sqrt(lo * hi) if hi / lo > 4 else 0.5 * (lo + hi)
        if f(mid) > 0:
            hi = mid
        else:
            lo = mid
    sigma = max(0.5 * (lo + hi), 1e-3 * d.mean())
    out = np.ones_like(d)
    out[act] = np.exp(-(d[act] - rho) / sigma)
    return out


def _graph(n, k, seed, scale=1.0):
    rng = np.random.default_rng(seed)
    dist = np.sort(rng.gamma(2.0, scale, size=(n, k)).astype(np.float32), axis=1)
    dist[:, 0] = 0.0
    ids = np.stack([np.concatenate([[i], rng.choice(10 * n, size=k - 1, replace=False)]) for i in range(n)]).astype(np.int64)
    return ids, dist


@pytest.mark.parametrize("k", [16, 91, 301])
def test_rows_satisfy_the_defining_equation(oracle, k):
    ids, dist = _graph(64, k, seed=k)
    cols, vals = oracle.prow_exponential_graph(ids, dist)
    np.testing.assert_array_equal(cols, ids[:, 1:].astype(np.uint32))  # neighbour order kept (MapMemEff instantiation)
    # the nearest real neighbour gets weight 1 (d == rho), everything lies in (0, 1], rows are non-increasing like the distances
    assert (vals[:, 0] == 1.0).all() and (vals > 0).all() and (vals <= 1.0).all()
    assert (np.diff(vals, axis=1) <= 0).all()
    np.testing.assert_allclose(vals.sum(axis=1), np.log2(k), atol=2e-5 * k)
    for p in range(0, 64, 7):
        np.testing.assert_allclose(vals[p], _numpy_row(dist[p, 1:]), rtol=2e-4, atol=1e-7)


def test_sorted_instantiation_is_a_permutation(oracle):
    ids, dist = _graph(32, 40, seed=5)
    c0, v0 = oracle.prow_exponential_graph(ids, dist, sort_cols=False)
    c1, v1 = oracle.prow_exponential_graph(ids, dist, sort_cols=True)
    assert (np.diff(c1.astype(np.int64), axis=1) > 0).all()
    for p in range(32):
        o = np.argsort(c0[p], kind="stable")
        np.testing.assert_array_equal(c0[p][o], c1[p])
        np.testing.assert_array_equal(v0[p][o], v1[p])


def test_edge_rows(oracle):
    k = 9
    ids = np.tile(np.arange(k, dtype=np.int64), (5, 1))
    dist = np.zeros((5, k), np.float32)
    dist[1, 1:] = 2.5                                   # all neighbours at the same distance: all <= rho -> weights 1
    dist[2, 1:] = [0, 1, 2, 3, 4, 5, 6, 7]              # a leading zero distance: rho = first NON-zero distance
    dist[3, 1:] = [1, 1, 1, 1, 1, 1, 1, 1e6]            # one far outlier
    dist[4, 1:] = [1e-3, 1, 1, 1, 1, 1, 1, 1]           # 7 weights would have to reach 2.17/7 each
    _, vals = oracle.prow_exponential_graph(ids, dist)
    assert (vals[0] == 1).all() and (vals[1] == 1).all()
    assert (vals[2, :2] == 1).all() and (vals[2, 2:] < 1).all() and (vals[2, 2:] > 0).all()
    np.testing.assert_allclose(vals[2], _numpy_row(dist[2, 1:]), rtol=2e-4, atol=1e-7)
    # 7 entries at rho give 7 > log2(9) = 3.17 already: the sum can only approach 7 from above, sigma shrinks to the bisection
    # floor and is then lifted by the clamp 1e-3 * mean(d) = 125.0 -> exp(-(1e6 - 1) / 125) underflows to 0
    assert (vals[3, :7] == 1).all() and vals[3, 7] == 0.0
    np.testing.assert_allclose(vals[4], _numpy_row(dist[4, 1:]), rtol=2e-4, atol=1e-7)


def test_inplace_rows_and_ignored_column(oracle):
    rng = np.random.default_rng(11)
    lens = rng.integers(0, 60, size=50)
    rowptr = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    col = np.concatenate([np.sort(rng.choice(200, size=m, replace=False)) for m in lens]).astype(np.uint32)
    val = rng.gamma(2.0, 1.0, size=rowptr[-1]).astype(np.float32)   # NOT sorted by value: stored (column) order counts
    out = oracle.prow_exponential_inplace(rowptr, val)
    for p in range(50):
        seg = slice(rowptr[p], rowptr[p + 1])
        np.testing.assert_allclose(out[seg], _numpy_row(val[seg]) if lens[p] else [], rtol=2e-4, atol=1e-7)
    out7 = oracle.prow_exponential_inplace(rowptr, val, col=col, ignore=7)
    for p in range(50):
        seg = slice(rowptr[p], rowptr[p + 1])
        keep = col[seg] != 7
        np.testing.assert_array_equal(out7[seg][~keep], val[seg][~keep])     # the ignored entry keeps its value
        if keep.any():
            np.testing.assert_allclose(out7[seg][keep], _numpy_row(val[seg][keep]), rtol=2e-4, atol=1e-7)
