"""GPU parity tests on the BASELINE.json configurations themselves: the kernel flavours, the
1-pass band, the pilot / auto switch, the overflow re-sweep and the SIMT fallback run here exactly as in bench.py.
Sampled query rows of the CUDA result (ids, fp32 distances, level-0 P rows) are recomputed from scratch by the oracle
(orc_knn_exact_rows + orc_knn_postprocess_rows + orc_prow_gaussian_graph): ids and distances bit-exact, P-row columns
bit-exact, P-row values rtol 1e-5 (north_star: "distances and similarities within 1e-5 relative").
"""
import numpy as np
import pytest

from sph_b200 import synth

pytestmark = pytest.mark.gpu
K = 91


def _sample_rows(capi, q0, q1, count, seed):
    rng = np.random.default_rng(seed)
    rows = q0 + rng.choice(q1 - q0, size=min(count, q1 - q0), replace=False)
    extra = [capi.knn_last_overflow_rows(0)[:24], capi.knn_last_overflow_rows(1)[:24]]
    return np.unique(np.concatenate([rows] + extra).astype(np.int64))


def _check_slab(capi, oracle, x, q0, q1, count, seed, expect_tensor=True):
    ids, dist, cols, vals = capi.knn_graph_prow(x, K, 0.0, q0=q0, q1=q1)
    st = capi.knn_last_stats()
    assert (st["tensor_core"] >= 1) == expect_tensor, st
    rows = _sample_rows(capi, q0, q1, count, seed)
    sel = rows - q0
    rids, rdist = oracle.knn_reference_graph(x, K, rows=rows)
    bad = np.nonzero((ids[sel] != rids).any(axis=1))[0]
    assert bad.size == 0, f"{bad.size} of {len(rows)} rows differ; first: row {rows[bad[0]]} gpu {ids[sel][bad[0]]} ref {rids[bad[0]]}"
    np.testing.assert_array_equal(dist[sel], rdist)
    rcols, rvals, _ = oracle.prow_gaussian_graph(rids, rdist, 0.0)
    np.testing.assert_array_equal(cols[sel], rcols)
    np.testing.assert_allclose(vals[sel], rvals, rtol=1e-5, atol=1e-12)
    # structural invariants of NearestNeighbors::compute over the WHOLE slab (sph/NearestNeighbors.cpp:143-170)
    assert (ids[:, 0] == np.arange(q0, q1)).all()
    assert (np.diff(dist, axis=1) >= 0).all()
    assert (ids >= 0).all() and (ids < x.shape[0]).all()
    return st, rows


def test_c1_full(capi, oracle):
    x = synth.make("C1")
    st, _ = _check_slab(capi, oracle, x, 0, x.shape[0], 512, 1)
    assert st["passes"] == 1


def test_c2_full(capi, oracle):
    x = synth.make("C2")
    _check_slab(capi, oracle, x, 0, x.shape[0], 256, 2)


def test_c3_full_width_strip(capi, oracle):
    """1920 x 64 strip of the RGB configuration: d = 3 -> exact SIMT kernel; quantised colours give exact duplicates and ties."""
    x = synth.rgb(64, 1920, seed=1003)
    assert len(np.unique(x, axis=0)) < x.shape[0]  # duplicates present
    _check_slab(capi, oracle, x, 0, x.shape[0], 256, 3, expect_tensor=False)


def test_c4_slab(capi, oracle):
    """A slab of the 2048 x 2048 x 128 stack against the FULL 4.19 M-row database: more than 3 waves of query tiles, so the pilot
    wave, the main 1-pass sweep (lock-step pacing, 32 768 database tiles), the overflow re-sweep and the re-rank all run."""
    x = synth.make("C4")
    n = x.shape[0]
    q0 = 1_000_000
    q1 = q0 + 148 * 256 * 3 + 4000
    st, rows = _check_slab(capi, oracle, x, q0, q1, 96, 4)
    assert st["tensor_core"] == 3 and st["passes"] == 1 and st["kp"] == 144
    assert n == 2048 * 2048


def test_c5_slab(capi, oracle):
    """A slab of the 1024 x 1024 x 1000 stack: d = 1000 -> streaming tcgen05 kernel, K extent ~ 1008 (1 pass) / 3008 (3 passes)."""
    x = synth.make("C5")
    q0 = 500_000
    q1 = q0 + 148 * 128 + 500
    st, _ = _check_slab(capi, oracle, x, q0, q1, 48, 5)
    assert st["tensor_core"] == 1
