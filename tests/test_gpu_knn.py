"""GPU parity tests, Stage A: CUDA exact kNN (through the C ABI) vs the CPU oracle.

Bar: bit-exact neighbour ids (ties broken by index) and bit-exact fp32 distances —
both sides evaluate the same fp64 FMA chain and round once.
"""
import numpy as np
import pytest

from sph_b200 import synth

pytestmark = pytest.mark.gpu


def _check(capi, oracle, x, k, path, l2_squared=False, rows=None, metric=1):
    capi.force_knn_path(path)
    try:
        ids, dist = capi.knn_exact(x, k, metric=metric, l2_squared=l2_squared)
    finally:
        capi.force_knn_path(0)
    if rows is None:
        rids, rdist = oracle.knn_reference_graph(x, k, l2_squared=l2_squared, metric=metric)
        gids, gdist = ids, dist
    else:
        rids, rdist = oracle.knn_reference_graph(x, k, l2_squared=l2_squared, rows=rows, metric=metric)
        gids, gdist = ids[rows], dist[rows]
    bad = np.nonzero((gids != rids).any(axis=1))[0]
    assert bad.size == 0, f"{bad.size} rows differ; first {bad[:5]}: gpu {gids[bad[0]]} ref {rids[bad[0]]}"
    np.testing.assert_array_equal(gdist, rdist)
    return ids, dist


@pytest.mark.parametrize("n,d,k", [(300, 3, 10), (1000, 7, 33), (777, 40, 91)])
def test_simt_matches_oracle(capi, oracle, n, d, k):
    rng = np.random.default_rng(n + d)
    x = rng.standard_normal((n, d)).astype(np.float32)
    _check(capi, oracle, x, k, path=1)


def test_simt_scurve_k3(capi, oracle):
    # shape of the reference's testConnectingComponentsKnnGraph (tests/UtilsTests.cpp:662-703)
    x = synth.s_curve(1000, 0.2, 123)
    ids, dist = _check(capi, oracle, x, 3, path=1)
    assert (ids[:, 0] == np.arange(1000)).all()
    assert (np.diff(dist, axis=1) >= 0).all()


def test_simt_duplicates_and_ties(capi, oracle):
    # quantised RGB crop: exact duplicates, lattice ties, rows whose k nearest are all at distance 0
    x = synth.rgb(48, 64, seed=7)
    x[:200] = x[0]  # 200 identical pixels -> self not among the first k by index for most of them
    _check(capi, oracle, x, 16, path=1)
    _check(capi, oracle, x, 16, path=1, l2_squared=True)


# path 2: auto flavour (resident MH=2 for d<=128, MH=1 for d<=256, streaming above), 3: streaming, 4: resident MH=1
@pytest.mark.parametrize("n,d,k,path,flavour", [
    (2048, 32, 16, 2, 3), (4100, 128, 91, 2, 3), (3000, 200, 91, 2, 2), (2500, 21, 5, 2, 3), (2300, 300, 40, 2, 1),
    (4100, 128, 91, 3, 1), (2500, 21, 5, 3, 1), (3000, 100, 30, 4, 2), (2048, 64, 301, 2, 3),
])
def test_tc_matches_oracle(capi, oracle, n, d, k, path, flavour):
    rng = np.random.default_rng(n * 7 + d)
    x = (rng.standard_normal((n, d)) * 0.3 + rng.uniform(0, 1, (1, d))).astype(np.float32)
    _check(capi, oracle, x, k, path=path)
    st = capi.knn_last_stats()
    assert st["tensor_core"] == flavour


def test_tc_hyperspectral_c1_shape(capi, oracle):
    # Indian-Pines-shaped stack (C1), k = 91; oracle on a sample of rows
    x = synth.hyperspectral(145, 145, 200, endmembers=16, seed=1001)
    rows = np.random.default_rng(0).choice(x.shape[0], 256, replace=False)
    rows.sort()
    _check(capi, oracle, x, 91, path=2, rows=rows)


def test_tc_duplicates(capi, oracle):
    rng = np.random.default_rng(5)
    base = rng.standard_normal((600, 16)).astype(np.float32)
    x = np.concatenate([base, base, base[:300], np.round(base * 4) / 4])  # exact duplicates + a coarse lattice
    _check(capi, oracle, x, 8, path=2)
    _check(capi, oracle, x, 130, path=2)
    _check(capi, oracle, x, 8, path=3)
    # heavy ties: 64 copies of each of 40 points -> the candidate band overflows and rows fall back to the exact kernel
    y = np.repeat(rng.standard_normal((40, 16)).astype(np.float32), 64, axis=0)
    _check(capi, oracle, y, 100, path=2)
    assert capi.knn_last_stats()["overflow_rows"] >= 0


def test_tc_equals_simt_on_device(capi):
    x = synth.hyperspectral(96, 128, 64, endmembers=8, seed=3)
    capi.force_knn_path(1)
    a = capi.knn_exact(x, 31)
    capi.force_knn_path(2)
    b = capi.knn_exact(x, 31)
    capi.force_knn_path(0)
    np.testing.assert_array_equal(a[0], b[0])
    np.testing.assert_array_equal(a[1], b[1])


def test_slab_equals_full(capi):
    x = synth.hyperspectral(64, 80, 48, endmembers=8, seed=4)
    full = capi.knn_exact(x, 20)
    n = x.shape[0]
    cuts = [0, 1000, 1001, 3333, n]
    for a, b in zip(cuts[:-1], cuts[1:]):
        ids, dist = capi.knn_exact(x, 20, q0=a, q1=b)
        np.testing.assert_array_equal(ids, full[0][a:b])
        np.testing.assert_array_equal(dist, full[1][a:b])


def test_invalid_arguments(capi):
    x = np.zeros((10, 4), np.float32)
    with pytest.raises(capi.SphbError):
        capi.knn_exact(x, 0)
    with pytest.raises(capi.SphbError):
        capi.knn_exact(x, 3, q0=5, q1=20)
