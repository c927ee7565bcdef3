"""GPU parity tests of the resident level pipeline (csrc/level_api.cu): the `_dev` entry points and the fused
sphb_level_prob_dist (LevelSimilarities::computeProbDistOnLevel's useRandomWalks branch + symmetrizeOutput,
sph/LevelSimilarities.cpp:460-508, 566-581, 589-623) against the staged host-buffer calls and the CPU oracle; both
accumulator homes of the createSimilarities kernel (shared-memory DENSE rows, L2 SLAB)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def walk_like_rows(n, seed, lo=5, hi=120, window=150):
    """rows shaped like merged random walks: some tens of columns from a local window, row-stochastic"""
    rng = np.random.default_rng(seed)
    rowptr = [0]
    col, val = [], []
    for i in range(n):
        m = int(rng.integers(lo, hi))
        c = np.unique(np.clip(i + rng.integers(-window, window, size=m), 0, n - 1))
        w = rng.gamma(0.7, 1.0, len(c)) + 1e-6
        col.extend(c.tolist())
        val.extend((w / w.sum()).tolist())
        rowptr.append(len(col))
    return np.array(rowptr, np.int64), np.array(col, np.uint32), np.array(val, np.float32)


def staged_gpu(capi, rp, c, v, k, prune, variant, perplexity, pair_sims=True, comp_sizes=None):
    if pair_sims:
        sc, sv, snz = capi.create_similarities(rp, c, v, k, prune, variant, comp_sizes)
    else:
        sc, sv, snz = capi.find_k(rp, c, v, k, top=True)
    prp = np.concatenate([[0], np.cumsum(snz, dtype=np.int64)])
    keep = np.arange(sc.shape[1])[None, :] < snz[:, None]
    pc, pv = sc[keep], sv[keep]
    pv = capi.prow_gaussian_inplace(prp, pv, perplexity, -1)
    zrp, zc, zv = capi.remove_zeros(prp, pc, pv)
    return capi.symmetrize(zrp, zc, zv)


def staged_oracle(oracle, rp, c, v, k, prune, variant, perplexity):
    sc, sv, snz = oracle.create_similarities(rp, c, v, k, prune, variant)
    prp = np.concatenate([[0], np.cumsum(snz, dtype=np.int64)])
    keep = np.arange(sc.shape[1])[None, :] < snz[:, None]
    pc, pv = sc[keep], sv[keep]
    pv = oracle.prow_gaussian_inplace(prp, pv, perplexity, -1)
    zrp, zc, zv = oracle.remove_zeros(prp, pc, pv)
    return oracle.symmetrize(zrp, zc, zv)


@pytest.mark.parametrize("variant,k,pair_sims", [(1, 31, True), (2, 31, True), (0, 91, True), (0, 16, False)])
def test_fused_level_equals_staged(capi, variant, k, pair_sims):
    rp, c, v = walk_like_rows(1500, seed=100 + k)
    perplexity = 10.0
    a = capi.level_prob_dist(rp, c, v, k, 1e-4, variant, None, pair_sims, perplexity, 1, True)
    b = staged_gpu(capi, rp, c, v, k, 1e-4, variant, perplexity, pair_sims)
    for x, y in zip(a, b):
        np.testing.assert_array_equal(x, y)   # same kernels, deterministic accumulation: bit-identical
    t = capi.level_last_timing()
    assert t["similarities_ms"] > 0 and t["symmetrize_ms"] > 0


def test_fused_level_vs_oracle(capi, oracle):
    rp, c, v = walk_like_rows(1200, seed=7)
    k, perplexity = 31, 10.0
    grp, gc, gv = capi.level_prob_dist(rp, c, v, k, 1e-4, 1, None, True, perplexity, 1, True)
    orp, oc, ov = staged_oracle(oracle, rp, c, v, k, 1e-4, 1, perplexity)
    n = len(rp) - 1
    same_rows = 0
    for i in range(n):
        a, b = slice(grp[i], grp[i + 1]), slice(orp[i], orp[i + 1])
        if np.array_equal(gc[a], oc[b]):
            np.testing.assert_allclose(gv[a], ov[b], rtol=2e-4, atol=1e-7)
            same_rows += 1
    # a 1-ulp difference of two similarities at the k-th slot of one row changes that row's pattern (and, through the
    # symmetrisation, one entry of the partner row); everything else must agree
    assert same_rows >= n - max(4, n // 100), same_rows


def test_fused_level_no_symmetrize_and_sizes(capi, oracle):
    rp, c, v = walk_like_rows(800, seed=3)
    rng = np.random.default_rng(4)
    sizes = rng.integers(1, 40, size=800)
    k = 20
    grp, gc, gv = capi.level_prob_dist(rp, c, v, k, 1e-4, 1, sizes, True, 0.0, 1, False)
    sc, sv, snz = capi.create_similarities(rp, c, v, k, 1e-4, 1, sizes)
    prp = np.concatenate([[0], np.cumsum(snz, dtype=np.int64)])
    keep = np.arange(sc.shape[1])[None, :] < snz[:, None]
    pv = capi.prow_gaussian_inplace(prp, sv[keep], 0.0, -1)
    zrp, zc, zv = capi.remove_zeros(prp, sc[keep], pv)
    np.testing.assert_array_equal(grp, zrp)
    np.testing.assert_array_equal(gc, zc)
    np.testing.assert_array_equal(gv, zv)


def _sims_equal_up_to_ties(gc, gv, gn, oc, ov, on):
    n = len(gn)
    bad = 0
    for i in range(n):
        if gn[i] != on[i] or not np.array_equal(gc[i, :gn[i]], oc[i, :on[i]]):
            assert gn[i] == on[i]
            assert len(set(gc[i, :gn[i]].tolist()) ^ set(oc[i, :on[i]].tolist())) <= 2
            bad += 1
            continue
        np.testing.assert_allclose(gv[i, :gn[i]], ov[i, :on[i]], rtol=2e-5, atol=2e-6)
    assert bad <= max(1, n // 200)


@pytest.mark.parametrize("variant", [1, 2])
def test_similarities_slab_and_dense_agree(capi, oracle, variant):
    """the shared-memory row (default when it fits) and the L2-slab kernel (large levels) compute the same rows"""
    rp, c, v = walk_like_rows(1500, seed=31 + variant)
    dense = capi.create_similarities(rp, c, v, 31, 1e-4, variant)
    capi.force_sims_slab(True)
    try:
        slab = capi.create_similarities(rp, c, v, 31, 1e-4, variant)
    finally:
        capi.force_sims_slab(False)
    ref = oracle.create_similarities(rp, c, v, 31, 1e-4, variant)
    _sims_equal_up_to_ties(*dense, *ref)
    _sims_equal_up_to_ties(*slab, *ref)
    again = capi.create_similarities(rp, c, v, 31, 1e-4, variant)
    for x, y in zip(dense, again):
        np.testing.assert_array_equal(x, y)   # the DENSE kernel accumulates in a fixed order


def test_similarities_dense_long_columns(capi, oracle):
    """a hub column present in every row (column length n > the 256-entry segment) and rows longer than one entry batch"""
    rng = np.random.default_rng(77)
    n = 2500
    rowptr, col, val = [0], [], []
    for i in range(n):
        m = 700 if i % 500 == 0 else int(rng.integers(4, 40))
        cset = np.unique(np.concatenate([[7], rng.integers(0, n, size=m)]))
        w = rng.uniform(0.01, 1.0, len(cset))
        col.extend(cset.tolist())
        val.extend((w / w.sum()).tolist())
        rowptr.append(len(col))
    rp, c, v = np.array(rowptr, np.int64), np.array(col, np.uint32), np.array(val, np.float32)
    g = capi.create_similarities(rp, c, v, 25, 1e-4, 1)
    o = oracle.create_similarities(rp, c, v, 25, 1e-4, 1)
    _sims_equal_up_to_ties(*g, *o)


def test_dev_entry_points_match_host(capi):
    """every `_dev` export against its host-buffer twin on device-resident copies (torch tensors as buffers)"""
    import torch

    rp, c, v = walk_like_rows(1000, seed=55)
    n, nnz = len(rp) - 1, len(c)
    s = torch.cuda.current_stream().cuda_stream
    drp = torch.from_numpy(rp).cuda()
    dc = torch.from_numpy(c.view(np.int32)).cuda()
    dv = torch.from_numpy(v).cuda()
    # rw_merge
    parents = (np.arange(n) // 3).astype(np.uint64)
    nm = int(parents.max()) + 1
    dpar = torch.from_numpy(parents.view(np.int64)).cuda()
    orp = torch.empty(nm + 1, dtype=torch.int64, device="cuda")
    oc = torch.empty(nnz, dtype=torch.int32, device="cuda")
    ov = torch.empty(nnz, dtype=torch.float32, device="cuda")
    m = capi.rw_merge_dev(n, drp.data_ptr(), dc.data_ptr(), dv.data_ptr(), nnz, dpar.data_ptr(), nm, True, True, orp.data_ptr(), oc.data_ptr(),
                          ov.data_ptr(), nnz, s)
    hrp, hc, hv = capi.rw_merge(rp, c, v, parents, nm, norm=True, weight_by_size=True)
    assert m == len(hc)
    np.testing.assert_array_equal(orp.cpu().numpy(), hrp)
    np.testing.assert_array_equal(oc[:m].cpu().numpy().view(np.uint32), hc)
    np.testing.assert_array_equal(ov[:m].cpu().numpy(), hv)
    # random walks (slab), create_similarities, in-place rows, symmetrise, find_k
    W, L = 20, 5
    cap = n * W * L
    wrp = torch.empty(n + 1, dtype=torch.int64, device="cuda")
    wc = torch.empty(cap, dtype=torch.int32, device="cuda")
    wv = torch.empty(cap, dtype=torch.float32, device="cuda")
    wm = capi.random_walks_dev(n, drp.data_ptr(), dc.data_ptr(), dv.data_ptr(), W, L, capi.W_NORMAL, 9, 0, n, wrp.data_ptr(), wc.data_ptr(),
                               wv.data_ptr(), cap, s)
    h = capi.random_walks(rp, c, v, W, L, capi.W_NORMAL, seed=9)
    assert wm == len(h[1])
    np.testing.assert_array_equal(wrp.cpu().numpy(), h[0])
    np.testing.assert_array_equal(wc[:wm].cpu().numpy().view(np.uint32), h[1])
    np.testing.assert_array_equal(wv[:wm].cpu().numpy(), h[2])
    k = 17
    sc = torch.empty((n, k), dtype=torch.int32, device="cuda")
    sv = torch.empty((n, k), dtype=torch.float32, device="cuda")
    sn = torch.empty(n, dtype=torch.int32, device="cuda")
    capi.create_similarities_dev(n, drp.data_ptr(), dc.data_ptr(), dv.data_ptr(), nnz, k, 1e-4, 1, None, sc.data_ptr(), sv.data_ptr(), sn.data_ptr(), s)
    hc2, hv2, hn2 = capi.create_similarities(rp, c, v, k, 1e-4, 1)
    np.testing.assert_array_equal(sn.cpu().numpy(), hn2)
    np.testing.assert_array_equal(sc.cpu().numpy().view(np.uint32), hc2)
    np.testing.assert_array_equal(sv.cpu().numpy(), hv2)
    fc = torch.empty((n, k), dtype=torch.int32, device="cuda")
    fv = torch.empty((n, k), dtype=torch.float32, device="cuda")
    fn = torch.empty(n, dtype=torch.int32, device="cuda")
    capi.find_k_dev(n, drp.data_ptr(), dc.data_ptr(), dv.data_ptr(), 0, k, True, fc.data_ptr(), fv.data_ptr(), fn.data_ptr(), s)
    hfc, hfv, hfn = capi.find_k(rp, c, v, k, top=True)
    np.testing.assert_array_equal(fn.cpu().numpy(), hfn)
    np.testing.assert_array_equal(fv.cpu().numpy(), hfv)
    dv2 = dv.clone()
    capi.prow_gaussian_inplace_dev(n, drp.data_ptr(), dv2.data_ptr(), 0, 8.0, -1, s)
    np.testing.assert_array_equal(dv2.cpu().numpy(), capi.prow_gaussian_inplace(rp, v, 8.0, -1))
    yrp = torch.empty(n + 1, dtype=torch.int64, device="cuda")
    yc = torch.empty(2 * nnz, dtype=torch.int32, device="cuda")
    yv = torch.empty(2 * nnz, dtype=torch.float32, device="cuda")
    ym = capi.symmetrize_dev(n, drp.data_ptr(), dc.data_ptr(), dv.data_ptr(), nnz, False, yrp.data_ptr(), yc.data_ptr(), yv.data_ptr(), 2 * nnz, s)
    hs = capi.symmetrize(rp, c, v)
    assert ym == len(hs[1])
    np.testing.assert_array_equal(yrp.cpu().numpy(), hs[0])
    np.testing.assert_array_equal(yv[:ym].cpu().numpy(), hs[2])
    # fused, device resident
    prp_ = torch.empty(n + 1, dtype=torch.int64, device="cuda")
    pc_ = torch.empty(2 * n * k, dtype=torch.int32, device="cuda")
    pv_ = torch.empty(2 * n * k, dtype=torch.float32, device="cuda")
    pm = capi.level_prob_dist_dev(n, drp.data_ptr(), dc.data_ptr(), dv.data_ptr(), nnz, k, 1e-4, 1, None, True, 5.0, 1, True, prp_.data_ptr(),
                                  pc_.data_ptr(), pv_.data_ptr(), 2 * n * k, s)
    hp = capi.level_prob_dist(rp, c, v, k, 1e-4, 1, None, True, 5.0, 1, True)
    assert pm == len(hp[1])
    np.testing.assert_array_equal(prp_.cpu().numpy(), hp[0])
    np.testing.assert_array_equal(pc_[:pm].cpu().numpy().view(np.uint32), hp[1])
    np.testing.assert_array_equal(pv_[:pm].cpu().numpy(), hp[2])


@pytest.mark.parametrize("norm", [1, 2])
@pytest.mark.parametrize("remove", [True, False])
def test_rw_post_merge_matches_host_restatement(capi, norm, remove):
    """RW4 after the merge (Hierarchy::updateRandomWalks, sph/utils/Hierarchy.cpp:287-307): diagonal + exact-zero removal with
    keepSingleEntry, ONEDIM / TWODIM normalisation in the reference's fp32 summation order — against the literal host loop."""
    from sph_b200 import host

    rp, c, v = walk_like_rows(700, seed=91, lo=1, hi=40, window=30)
    v = v.copy()
    v[::17] = 0.0                      # stored zeros (removed together with the diagonal in rows with > 1 entries)
    rows = host.SparseRows(rp, c, v)
    want = host.removeDiagonalAndNormalize(rows, remove, host.NormType.ONEDIM if norm == 1 else host.NormType.TWODIM)
    got = capi.rw_post_merge(rp, c, v, remove, norm)
    np.testing.assert_array_equal(got[0], want.rowptr)
    np.testing.assert_array_equal(got[1], want.col)
    np.testing.assert_array_equal(got[2], want.val)
