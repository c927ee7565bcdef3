"""GPU tests of the host merge loop (sph_b200.host.ImageHierarchy) running on the CUDA path through the C ABI: the
reference's full-hierarchy goldens (tests/HierarchyTest.cpp:230-361, 635-784) must come out of the device kernels exactly
as they come out of the CPU oracle, and the random-walk hierarchy (Philox walks instead of the reference's mt19937_64)
must keep the structural invariants the reference asserts (row sums, nesting, reduction)."""
import json
from pathlib import Path

import numpy as np
import pytest

import oracle_lib
from test_oracle_goldens import check_hierarchy, run_hierarchy, run_walks_4x4

pytestmark = pytest.mark.gpu
GOLD = Path(__file__).resolve().parent / "golden"


def gold(name):
    return json.loads((GOLD / name).read_text())


@pytest.mark.parametrize("name", ["hierarchy_overlap_4x4.json", "hierarchy_overlap_6x4.json"])
def test_overlap_hierarchy_golden_on_device(capi, name):
    g = gold(name)
    ih = run_hierarchy(g, None)  # ops=None -> host.DeviceOps, the C ABI
    check_hierarchy(ih, g)


def test_overlap_hierarchy_matches_oracle_on_synthetic_image(capi, oracle):
    # a 24 x 20 synthetic stack: real kNN graph from the device, hierarchy on the device vs the same loop on the oracle
    from sph_b200 import host, synth

    x = synth.hyperspectral(24, 20, 16, endmembers=4, seed=3)
    ids, dist = oracle.knn_reference_graph(x, 9)  # input only: the device kNN is bit-identical to it (test_gpu_knn.py)
    g = host.Graph(ids, dist)
    ihs = host.ImageHierarchySettings(componentSim=host.ComponentSim.NEIGH_OVERLAP, usePercentile=False)
    a = host.ImageHierarchy(g, x, 24, 20)
    a.compute(ihs)
    b = host.ImageHierarchy(g, x, 24, 20, ops=oracle_lib.OracleOps(oracle))
    b.compute(ihs)
    assert a.getNumLevels() == b.getNumLevels() >= 3
    for lvl in range(a.getNumLevels()):
        np.testing.assert_array_equal(a.pixelComponentsOn(lvl), b.pixelComponentsOn(lvl))


def test_walk_hierarchy_invariants_on_device(capi):
    g = gold("hierarchy_walks_6x4.json")
    ih = run_hierarchy(g, None)
    levels = ih.getNumLevels()
    assert levels >= 2 and ih.numComponents[-1] == 1  # thresh = FLT_MAX: every component merges with a spatial neighbour
    assert all(b < a for a, b in zip(ih.numComponents[:-1], ih.numComponents[1:]))
    for lvl in range(levels - 1):  # nesting: a pixel's component on lvl + 1 is the parent of its component on lvl
        np.testing.assert_array_equal(ih.pixelComponentsOn(lvl + 1), ih.parentsOn(lvl)[ih.pixelComponentsOn(lvl)].astype(np.int64))
    for rw in ih.randomWalks:  # testImageHierarchyWalks (:440-444)
        sums = np.add.reduceat(rw.val, rw.rowptr[:-1])
        np.testing.assert_allclose(sums, 1.0, atol=1e-3)


def test_image_hierarchy_walks_row_sums_on_device(capi):
    run_walks_4x4(None)  # tests/HierarchyTest.cpp:364-445 through the C ABI
