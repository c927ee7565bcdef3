"""SURVEY 8(e), createSimilarities sharded by OUTPUT ROW: sphb_create_similarities_slab[_dev] (rows [r0, r1) of the result against
the whole random-walk matrix) and sphb_level_from_sims_dev (the level pipeline behind the gathered slabs).  An output row needs
the whole matrix and nothing of the other output rows, so the bar is: slabs concatenated == the one-call result — bit for bit
for the shared-memory class flavour of the kernel (deterministic sums), to float-atomic order for the L2-slab flavour — and
slabs -> level tail == sphb_level_prob_dist."""
import os

import numpy as np
import pytest

from test_gpu_level import _sims_equal_up_to_ties, walk_like_rows

pytestmark = pytest.mark.gpu
EMU = os.environ.get("SPHB_EMULATE") == "1"


class Dev:
    """Device buffers for the `_dev` entry points: torch CUDA tensors on a B200; plain numpy arrays under the CPU emulation of
    tests/emu (where "device" memory is host memory)."""

    def __init__(self):
        if not EMU:
            import torch

            self.torch = torch
            self.stream = torch.cuda.current_stream().cuda_stream
        else:
            self.stream = 0

    def put(self, a):
        a = np.ascontiguousarray(a)
        if EMU:
            return a.copy()
        view = a.view(np.int32) if a.dtype == np.uint32 else a
        return self.torch.from_numpy(view.copy()).cuda()

    def empty(self, shape, dtype):
        if EMU:
            return np.zeros(shape, dtype)
        t = {np.int32: self.torch.int32, np.uint32: self.torch.int32, np.float32: self.torch.float32, np.int64: self.torch.int64}[dtype]
        return self.torch.zeros(shape, dtype=t, device="cuda")

    @staticmethod
    def ptr(b, offset_elems=0):
        if isinstance(b, np.ndarray):
            return b.ctypes.data + offset_elems * b.itemsize
        return b.data_ptr() + offset_elems * b.element_size()

    def get(self, b, dtype=None):
        if not isinstance(b, np.ndarray):
            self.torch.cuda.synchronize()
            b = b.cpu().numpy()
        return b.view(dtype) if dtype is not None else b


N_ROWS = 240 if EMU else 1200  # the emulation switches fibers at every warp collective: keep its matrices small
SLABS = [(0, 0), (0, 1), (1, 37), (37, 38), (38, 140), (140, 140), (140, N_ROWS - 1), (N_ROWS - 1, N_ROWS)]


@pytest.mark.parametrize("variant,k,force_slab", [(1, 31, False), (2, 17, False), (1, 31, True), (0, 91, False)])
def test_row_slabs_concatenate_to_the_one_call_result(capi, variant, k, force_slab):
    rp, c, v = walk_like_rows(N_ROWS, seed=7 + k, hi=120 if not EMU else 40)
    sizes = np.random.default_rng(k).integers(1, 40, N_ROWS) if variant == 1 else None
    capi.force_sims_slab(force_slab)
    try:
        full = capi.create_similarities(rp, c, v, k, 1e-4, variant, sizes)
        parts = [capi.create_similarities_slab(rp, c, v, k, r0, r1, 1e-4, variant, sizes) for r0, r1 in SLABS]
    finally:
        capi.force_sims_slab(False)
    for (r0, r1), p in zip(SLABS, parts):
        assert p[0].shape == (r1 - r0, k) and len(p[2]) == r1 - r0
    cat = [np.concatenate([p[j] for p in parts]) for j in range(3)]
    if force_slab:
        # the L2-slab flavour adds its products with float atomics, whose order is not fixed: two calls on the same rows agree to
        # the last bits only, and a k-th-slot tie may swap (the comparison of test_similarities_slab_and_dense_agree)
        _sims_equal_up_to_ties(cat[0], cat[1], cat[2], *full)
        return
    np.testing.assert_array_equal(cat[2], full[2])  # shared-memory class flavour: every sum in ascending column order, deterministic
    np.testing.assert_array_equal(cat[0], full[0])
    np.testing.assert_array_equal(cat[1], full[1])


def test_slab_argument_checks(capi):
    from sph_b200.capi import SphbError

    rp, c, v = walk_like_rows(50, seed=1, lo=2, hi=8, window=10)
    for r0, r1 in ((-1, 5), (10, 5), (0, 51)):
        with pytest.raises(SphbError):
            capi.create_similarities_slab(rp, c, v, 5, r0, r1)


@pytest.mark.parametrize("world", [1, 3])
def test_gathered_slabs_through_the_level_tail_equal_the_fused_level(capi, world):
    """What multigpu.ShardedLevel does on `world` devices, replayed on one: slab r of every "rank" written to its place in the
    gathered fixed-stride buffers (rank * chunk), then sphb_level_from_sims_dev."""
    from sph_b200.multigpu import slab_bounds

    dev = Dev()
    n, k, perplexity = (1000 if not EMU else 200), 23, 7.0
    rp, c, v = walk_like_rows(n, seed=77, hi=120 if not EMU else 40)
    nnz = len(c)
    drp, dc, dv = dev.put(rp), dev.put(c), dev.put(v)
    chunk = slab_bounds(n, 0, world)[2]
    cols = dev.put(np.full((world * chunk, k), 0xFFFFFFFF, np.uint32))
    vals = dev.empty((world * chunk, k), np.float32)
    rnz = dev.empty((world * chunk,), np.int32)
    for rank in range(world):
        r0, r1, _ = slab_bounds(n, rank, world)
        lo = rank * chunk
        assert lo == r0
        capi.create_similarities_slab_dev(n, dev.ptr(drp), dev.ptr(dc), dev.ptr(dv), nnz, k, 1e-4, 1, None, r0, r1, dev.ptr(cols, lo * k), dev.ptr(vals, lo * k),
                                          dev.ptr(rnz, lo), dev.stream)
    cap = 2 * n * k
    orp, oc, ov = dev.empty((n + 1,), np.int64), dev.empty((cap,), np.int32), dev.empty((cap,), np.float32)
    m = capi.level_from_sims_dev(n, k, dev.ptr(cols), dev.ptr(vals), dev.ptr(rnz), perplexity, 1, True, dev.ptr(orp), dev.ptr(oc), dev.ptr(ov), cap, dev.stream)
    hrp, hc, hv = capi.level_prob_dist(rp, c, v, k, 1e-4, 1, None, True, perplexity, 1, True)
    assert m == len(hc)
    np.testing.assert_array_equal(dev.get(orp), hrp)
    np.testing.assert_array_equal(dev.get(oc)[:m].view(np.uint32), hc)
    np.testing.assert_array_equal(dev.get(ov)[:m], hv)


@pytest.mark.skipif(EMU, reason="torch CUDA tensors")
def test_sharded_level_single_rank_equals_fused_level(capi):
    import torch

    from sph_b200.multigpu import ShardedLevel

    n, k, perplexity = 900, 19, 6.0
    rp, c, v = walk_like_rows(n, seed=91)
    job = ShardedLevel(0, 1)
    orp, oc, ov = job(torch.from_numpy(rp).cuda(), torch.from_numpy(c.view(np.int32)).cuda(), torch.from_numpy(v).cuda(), k, perplexity)
    torch.cuda.synchronize()
    hrp, hc, hv = capi.level_prob_dist(rp, c, v, k, 1e-4, 0, None, True, perplexity, 1, True)
    np.testing.assert_array_equal(orp.cpu().numpy(), hrp)
    np.testing.assert_array_equal(oc.cpu().numpy().view(np.uint32), hc)
    np.testing.assert_array_equal(ov.cpu().numpy(), hv)
