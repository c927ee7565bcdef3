"""GPU tests of the single-process multi-device entry point sphb_knn_graph_prow_multi (one worker thread per listed device, slabs
exchanged device to device): the result must be the bytes sphb_knn_graph_prow gives on one GPU.  On a one-GPU box the device list
names GPU 0 several times — the worker threads, the slab exchange (peer copies onto the same device) and the per-thread contexts
are exercised all the same; with two or more GPUs visible the list names distinct devices."""
import numpy as np
import pytest

from sph_b200 import synth

pytestmark = pytest.mark.gpu


def _devices(count):
    import torch

    have = torch.cuda.device_count()
    return [g % have for g in range(count)]


@pytest.mark.parametrize("workers,n_side,d,k", [(2, 96, 24, 31), (3, 75, 40, 16), (1, 64, 16, 10)])
def test_multi_equals_single(capi, workers, n_side, d, k):
    x = synth.hyperspectral(n_side, n_side, d, endmembers=5, seed=70 + workers)
    ref = capi.knn_graph_prow(x, k, 0.0)
    got = capi.knn_graph_prow_multi(x, k, _devices(workers))
    for a, b in zip(got, ref):
        np.testing.assert_array_equal(a, b)
    # the calling thread's own context is untouched and still works
    again = capi.knn_graph_prow(x, k, 0.0)
    for a, b in zip(again, ref):
        np.testing.assert_array_equal(a, b)


def test_multi_more_workers_than_rows_per_chunk(capi):
    x = np.random.default_rng(3).standard_normal((1500, 12)).astype(np.float32)
    ref = capi.knn_graph_prow(x, 8, 0.0)
    got = capi.knn_graph_prow_multi(x, 8, _devices(4))
    for a, b in zip(got, ref):
        np.testing.assert_array_equal(a, b)


def test_multi_reports_errors(capi):
    x = np.zeros((100, 4), np.float32)
    with pytest.raises(Exception, match="device"):
        capi.knn_graph_prow_multi(x, 5, [0, 99])


def test_sharded_call_overlapped_download(capi):
    """multigpu.ShardedKnnProw.__call__ (the N > 1 bench path) on one rank: the graph is copied back on a side stream while the P-row
    kernel runs; the host buffers must hold exactly what sphb_knn_graph_prow returns."""
    import torch

    from sph_b200 import multigpu

    x = synth.hyperspectral(80, 80, 32, endmembers=6, seed=11)
    n, d = x.shape
    k = 21
    ref = capi.knn_graph_prow(x, k, 0.0)
    job = multigpu.ShardedKnnProw(n, d, k, rank=0, world=1)
    out = (torch.empty((n, k), dtype=torch.int64).pin_memory(), torch.empty((n, k), dtype=torch.float32).pin_memory(),
           torch.empty((n, k - 1), dtype=torch.int32).pin_memory(), torch.empty((n, k - 1), dtype=torch.float32).pin_memory())
    xh = torch.from_numpy(x).pin_memory()
    for _ in range(2):
        for o in out:
            o.zero_()
        job(xh, out)
        torch.cuda.current_stream().synchronize()
        np.testing.assert_array_equal(out[0].numpy(), ref[0])
        np.testing.assert_array_equal(out[1].numpy(), ref[1])
        np.testing.assert_array_equal(out[2].numpy().view(np.uint32), ref[2])
        np.testing.assert_array_equal(out[3].numpy(), ref[3])
