"""GPU test of sphb_knn_rows_noalloc_dev (the no-allocation entry point added; no reference
counterpart): its output equals the existing sphb_knn_exact_dev's rows bit for bit, on two calls in a row (the fixed
scratch array keeps the previous call's contents), and sizes beyond the fixed scratch are rejected.  Collected last."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_noalloc_knn_rows_match_existing_entry_point(capi):
    import torch

    rng = np.random.default_rng(2024)
    n, d, k = 6000, 24, 31
    x = rng.standard_normal((n, d)).astype(np.float32)
    x[100:110] = x[100]  # duplicates: the self-first post-pass fires
    dx = torch.from_numpy(x).cuda()
    s = torch.cuda.current_stream().cuda_stream
    ids_ref = torch.empty((n, k), dtype=torch.int64, device="cuda")
    dist_ref = torch.empty((n, k), dtype=torch.float32, device="cuda")
    capi.knn_exact_dev(dx.data_ptr(), n, d, k, capi.METRIC_L2, False, 0, n, ids_ref.data_ptr(), dist_ref.data_ptr(), s)
    torch.cuda.synchronize()
    ids_ref, dist_ref = ids_ref.cpu().numpy(), dist_ref.cpu().numpy()
    for call, nrows in enumerate((256, 77)):  # the second call reuses the scratch left by the first, with another shape
        rows = rng.choice(n, size=nrows, replace=False).astype(np.int64)
        rows[:5] = [100, 101, 109, 0, n - 1]
        drows = torch.from_numpy(rows).cuda()
        ids = torch.full((nrows, k), -7, dtype=torch.int64, device="cuda")
        dist = torch.full((nrows, k), -7.0, dtype=torch.float32, device="cuda")
        free_before = torch.cuda.mem_get_info()[0]
        capi.knn_rows_noalloc_dev(dx.data_ptr(), n, d, k, capi.METRIC_L2, False, drows.data_ptr(), nrows, ids.data_ptr(), dist.data_ptr(), s)
        torch.cuda.synchronize()
        if call == 1:
            assert torch.cuda.mem_get_info()[0] >= free_before  # nothing was allocated on the device by the call
        np.testing.assert_array_equal(ids.cpu().numpy(), ids_ref[rows])
        np.testing.assert_array_equal(dist.cpu().numpy(), dist_ref[rows])


def test_noalloc_rejects_sizes_beyond_the_fixed_scratch(capi):
    import torch

    x = torch.zeros((2048, 8), dtype=torch.float32, device="cuda")
    rows = torch.zeros(300, dtype=torch.int64, device="cuda")
    ids = torch.empty((300, 200), dtype=torch.int64, device="cuda")
    dist = torch.empty((300, 200), dtype=torch.float32, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    with pytest.raises(capi.SphbError, match="256 query rows"):
        capi.knn_rows_noalloc_dev(x.data_ptr(), 2048, 8, 16, capi.METRIC_L2, False, rows.data_ptr(), 300, ids.data_ptr(), dist.data_ptr(), s)
    with pytest.raises(capi.SphbError, match="k > 128"):
        capi.knn_rows_noalloc_dev(x.data_ptr(), 2048, 8, 200, capi.METRIC_L2, False, rows.data_ptr(), 10, ids.data_ptr(), dist.data_ptr(), s)
