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


# paths (capi.force_knn_path): 2 auto (1-pass fp16 filter, 3-pass re-sweep of overflow rows), 3 streaming 3-pass,
# 4 resident MH=1 3-pass, 5 3-pass auto flavour, 6 1-pass, 7 1-pass streaming.
# flavour (stats["tensor_core"]): 1 streaming, 2 resident MH=1, 3 resident MH=2, 4 CTA pair (cta_group::2).
@pytest.mark.parametrize("n,d,k,path,flavour,passes", [
    (2048, 32, 16, 2, 3, 1), (4100, 128, 91, 2, 3, 1), (3000, 200, 91, 2, 3, 1), (2500, 21, 5, 2, 3, 1), (2300, 300, 40, 2, 2, 1),
    (2100, 600, 20, 2, 1, 1), (4100, 128, 91, 6, 3, 1), (2500, 21, 5, 7, 1, 1), (2048, 64, 301, 6, 3, 1), (40000, 256, 91, 6, 3, 1),
    (1100, 140, 20, 6, 3, 1), (3000, 100, 192, 6, 3, 1), (3000, 100, 193, 6, 3, 1),
    (4100, 128, 91, 5, 3, 3), (3000, 200, 91, 5, 2, 3), (2300, 300, 40, 5, 1, 3),
    (4100, 128, 91, 3, 1, 3), (2500, 21, 5, 3, 1, 3), (3000, 100, 30, 4, 2, 3), (2048, 64, 301, 5, 3, 3),
])
def test_tc_matches_oracle(capi, oracle, n, d, k, path, flavour, passes):
    rng = np.random.default_rng(n * 7 + d)
    x = (rng.standard_normal((n, d)) * 0.3 + rng.uniform(0, 1, (1, d))).astype(np.float32)
    _check(capi, oracle, x, k, path=path)
    st = capi.knn_last_stats()
    assert st["tensor_core"] == flavour
    assert st["passes"] == passes


def test_tc_norm_chunk_layout(capi, oracle):
    # the |v|² chunk is a narrow (16-column, SWIZZLE_32B) tile next to the 64-column data chunks, 1-pass and 3-pass sweeps
    x = synth.hyperspectral(64, 64, 96, endmembers=8, seed=21)
    _check(capi, oracle, x, 31, path=6)
    _check(capi, oracle, x, 31, path=5)


def test_tc_resweep_database_segments(capi, oracle):
    """Few overflow rows against a database of many tiles: the re-sweep cuts the database into segments (sweep_item), the
    re-rank merges the per-segment candidate lists; rows sit on both sides of segment borders."""
    rng = np.random.default_rng(5)
    n, d, k = 60000, 64, 40
    x = (rng.standard_normal((n, d)) * 0.3).astype(np.float32)
    # three tight clusters far from the mean, spread over the index range: their rows overflow the 1-pass band
    for c0 in (100, 29990, 59000):
        x[c0:c0 + 700] = (5.0 + rng.standard_normal((1, d)) + 1e-3 * rng.standard_normal((700, d))).astype(np.float32)
    rows = np.concatenate([np.arange(90, 140), np.arange(29980, 30050), np.arange(59600, 59710), rng.choice(n, 200, replace=False)])
    _check(capi, oracle, x, k, path=2, rows=np.unique(rows))
    st = capi.knn_last_stats()
    assert st["resweep_rows"] > 0
    assert len(capi.knn_last_overflow_rows(0)) == st["resweep_rows"]


def test_tc_one_pass_overflow_resweep(capi, oracle):
    # a tight cluster far from the data mean: |u| is large against the intra-cluster distances, the 1-pass
    # band holds the whole cluster -> those rows are re-swept with the 3-pass filter (and what still overflows
    # goes to the exact SIMT kernel); results stay exact
    rng = np.random.default_rng(17)
    a = (10.0 + 0.01 * rng.standard_normal((5000, 32))).astype(np.float32)
    b = (-50.0 + rng.standard_normal((1000, 32))).astype(np.float32)
    x = np.concatenate([a, b])
    rows = np.arange(0, 6000, 47)
    _check(capi, oracle, x, 91, path=6, rows=rows)
    st = capi.knn_last_stats()
    assert st["passes"] == 1 and st["resweep_rows"] > 0
    _check(capi, oracle, x, 91, path=2, rows=rows)


def test_tc_inner_product(capi, oracle):
    rng = np.random.default_rng(23)
    x = rng.standard_normal((3000, 48)).astype(np.float32)
    x /= np.linalg.norm(x, axis=1, keepdims=True)  # COSINE = renormalise + inner product (NearestNeighbors.cpp:208-222)
    for path in (1, 6, 5):
        _check(capi, oracle, x, 20, path=path, metric=0)


def test_tc_auto_pilot_large(capi, oracle):
    # more than three waves of query tiles: the auto path runs a one-wave 1-pass pilot first
    x = synth.hyperspectral(360, 360, 16, endmembers=6, seed=31)
    rows = np.random.default_rng(1).choice(x.shape[0], 96, replace=False)
    rows.sort()
    _check(capi, oracle, x, 12, path=0, rows=rows)


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
    _check(capi, oracle, x, 130, path=5)
    # heavy ties: 600 copies of each of 10 points -> the candidate band overflows in both filters and the rows
    # fall back to the exact SIMT kernel
    y = np.repeat(rng.standard_normal((10, 16)).astype(np.float32), 600, axis=0)
    _check(capi, oracle, y, 100, path=2)
    st = capi.knn_last_stats()
    assert st["resweep_rows"] > 0 and st["overflow_rows"] > 0
    _check(capi, oracle, y, 100, path=5)
    assert capi.knn_last_stats()["overflow_rows"] > 0


def test_tc_few_overflow_rows_segmented_fallback(capi, oracle):
    # 2 x 600 identical points inside 7000 ordinary ones: 1200 rows overflow both filters; so few rows go through the
    # database-segmented SIMT fallback (one warp per (row, segment) + merge)
    rng = np.random.default_rng(11)
    x = rng.standard_normal((8200, 24)).astype(np.float32)
    x[1000:1600] = x[1000]
    x[5000:5600] = x[5000]
    _check(capi, oracle, x, 100, path=2)
    st = capi.knn_last_stats()
    assert 0 < st["overflow_rows"] <= 2300
    _check(capi, oracle, x, 100, path=2, metric=0)


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


def test_cosine_metric_on_device(capi, oracle):
    """KnnMetric::COSINE: the back-end normalises a device copy of the rows (initFaissData, sph/NearestNeighbors.cpp:208-222) and
    searches by inner product; the normalised copy is bit-identical to the oracle's fvec_renorm_L2 restatement."""
    import ctypes as C

    rng = np.random.default_rng(77)
    x = (rng.standard_normal((3000, 40)) * rng.uniform(0.1, 10, (3000, 1))).astype(np.float32)
    x[17] = 0  # all-zero row: copied unchanged
    ref = np.empty_like(x)
    oracle.lib.orc_renorm_l2(x.ctypes.data_as(C.POINTER(C.c_float)), x.shape[0], x.shape[1], ref.ctypes.data_as(C.POINTER(C.c_float)))
    np.testing.assert_array_equal(capi.renorm_l2(x), ref)
    ids, dist = capi.knn_exact(x, 21, metric=capi.METRIC_COSINE)
    rids, rdist = oracle.knn_reference_graph(ref, 21, metric=0)
    np.testing.assert_array_equal(ids, rids)
    np.testing.assert_array_equal(dist, rdist)
