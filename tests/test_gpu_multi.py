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
