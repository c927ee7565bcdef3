"""GPU parity tests, SURVEY §8(f)-2: the device parts of NearestNeighbors::connectComponents (sph/NearestNeighbors.cpp:494-861)
through the C ABI against the oracle — centres of mass bit-exact (sequential fp64 sums), closest pairs bit-exact (ids and the
squared distance, ties to the lowest point of B, then of A) — and the reference's own test of the whole step
(tests/UtilsTests.cpp:662-703) with the device in the loop."""
import numpy as np
import pytest

from sph_b200 import host
from test_host_bridge import s_curve

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n,d,ncomp", [(5000, 3, 7), (3000, 128, 5), (700, 300, 40), (20000, 16, 3), (64, 5, 64)])
def test_component_centers_bit_exact(capi, oracle, n, d, ncomp):
    rng = np.random.default_rng(n + d)
    x = (rng.standard_normal((n, d)) * rng.gamma(2.0, 3.0, size=(n, 1))).astype(np.float32)
    labels = rng.integers(0, ncomp, size=n).astype(np.int64)
    labels[:ncomp] = np.arange(ncomp)          # no empty component
    got = capi.component_centers(x, labels, ncomp)
    ref = oracle.component_centers(x, labels, ncomp)
    np.testing.assert_array_equal(got, ref)


def _clusters(sizes, d, seed, spread=6.0, quant=None):
    rng = np.random.default_rng(seed)
    xs, ls = [], []
    for c, m in enumerate(sizes):
        centre = rng.standard_normal(d) * spread
        xs.append(centre + rng.standard_normal((m, d)))
        ls.append(np.full(m, c))
    x = np.concatenate(xs).astype(np.float32)
    labels = np.concatenate(ls).astype(np.int64)
    if quant:
        x = np.round(x * quant) / quant
    o = rng.permutation(len(x))                # components interleaved in memory, as connected components are
    return np.ascontiguousarray(x[o]), labels[o]


@pytest.mark.parametrize("d,quant", [(3, None), (3, 2.0), (16, None), (128, None), (130, None), (37, 1.0)])
def test_bridge_pairs_bit_exact(capi, oracle, d, quant):
    sizes = [1, 63, 64, 65, 1000, 2, 129, 4000, 300]
    x, labels = _clusters(sizes, d, seed=d, quant=quant)
    ncomp = len(sizes)
    edges = np.array([[a, b] for a in range(ncomp) for b in range(ncomp) if a != b], np.int64)   # both orientations of every pair
    ia, ib, dd = capi.bridge_components(x, labels, ncomp, edges)
    ra, rb, rd = oracle.bridge_components(x, labels, ncomp, edges)
    np.testing.assert_array_equal(dd, rd)
    np.testing.assert_array_equal(ib, rb)
    np.testing.assert_array_equal(ia, ra)
    assert (labels[ia] == edges[:, 0]).all() and (labels[ib] == edges[:, 1]).all()


def test_bridge_pairs_exact_ties(capi, oracle):
    """Duplicated points across components: many pairs at distance 0 -> lowest point of B, then lowest point of A."""
    rng = np.random.default_rng(4)
    base = rng.integers(0, 3, size=(40, 6)).astype(np.float32)
    x = np.concatenate([base, base, base + 0.5]).astype(np.float32)
    labels = np.concatenate([np.zeros(40), np.ones(40), np.full(40, 2)]).astype(np.int64)
    o = rng.permutation(120)
    x, labels = np.ascontiguousarray(x[o]), labels[o]
    edges = np.array([[0, 1], [1, 0], [0, 2], [2, 1]], np.int64)
    got = capi.bridge_components(x, labels, 3, edges)
    ref = oracle.bridge_components(x, labels, 3, edges)
    for g, r in zip(got, ref):
        np.testing.assert_array_equal(g, r)
    assert got[2][0] == 0.0 and got[2][1] == 0.0


def test_bridge_rejects_bad_arguments(capi):
    x = np.zeros((10, 3), np.float32)
    with pytest.raises(Exception, match="out of range"):
        capi.bridge_components(x, np.zeros(10, np.int64), 1, np.array([[0, 1]], np.int64))
    with pytest.raises(Exception, match="labels"):
        capi.bridge_components(x, np.full(10, 5, np.int64), 2, np.array([[0, 1]], np.int64))
    with pytest.raises(Exception, match="L2"):
        capi.bridge_components(x, np.zeros(10, np.int64), 1, np.array([[0, 0]], np.int64), metric=capi.METRIC_INNER_PRODUCT)


@pytest.mark.parametrize("symmetric", [False, True])
def test_connecting_components_on_the_device(capi, oracle, symmetric):
    """tests/UtilsTests.cpp:662-703 with every compute call on the GPU; the same flow with the checker injected gives the same graph."""
    x = s_curve(1000, 0.2, 123)
    nn = host.NearestNeighbors(x)
    nn.compute(host.NearestNeighborsSettings(numNearestNeighbors=3, symmetricNeighbors=symmetric, computeConnectComponents=True))
    num, labels = nn.getConnectedComponents()
    assert num > 1
    g = nn.connectComponents()
    assert g is not None
    new_num, _ = capi.weak_components(g.knnIndices, g.rowptr)
    assert new_num == 1
    ref = nn.connectComponents(bridge_ops=oracle)
    np.testing.assert_array_equal(g.rowptr, ref.rowptr)
    np.testing.assert_array_equal(g.knnIndices, ref.knnIndices)
    np.testing.assert_array_equal(g.knnDistances, ref.knnDistances)


def test_bridge_two_large_components(capi, oracle):
    """A pair of components far beyond the reference's exact-search limit (3900 points): the device result is the exact closest pair;
    checked against the oracle on the candidate rows only (the full 30 000 x 20 000 oracle scan would take minutes)."""
    rng = np.random.default_rng(8)
    d = 32
    a = rng.standard_normal((30000, d)).astype(np.float32)
    b = (rng.standard_normal((20000, d)) + 3.0).astype(np.float32)
    x = np.concatenate([a, b])
    labels = np.concatenate([np.zeros(30000), np.ones(20000)]).astype(np.int64)
    ia, ib, dd = capi.bridge_components(x, labels, 2, np.array([[0, 1]], np.int64))
    # exact reference by blocks in fp64 (numpy), then the contract distance of the winner
    best = (np.inf, -1, -1)
    a64 = a.astype(np.float64)
    for s in range(0, 20000, 500):
        blk = b[s:s + 500].astype(np.float64)
        dist = (blk * blk).sum(1)[:, None] + (a64 * a64).sum(1)[None, :] - 2.0 * blk @ a64.T
        j = np.unravel_index(np.argmin(dist), dist.shape)
        if dist[j] < best[0]:
            best = (dist[j], 30000 + s + j[0], j[1])
    assert (int(ia[0]), int(ib[0])) == (best[2], best[1])
    sub = np.concatenate([x[ia], x[ib]])
    _, _, rd = oracle.bridge_components(sub, np.array([0, 1], np.int64), 2, np.array([[0, 1]], np.int64))
    np.testing.assert_array_equal(dd, rd)
