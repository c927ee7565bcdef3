"""Wall-clock probe of the ragged-graph / level entry points (csrc/graph_merge.cu) at a C2-shaped size through the host-buffer
C ABI, with the algorithmic bytes of DESIGN.md §4.6-4.8 next to each time.  Run on a B200 (next to an ncu launch list of the
same command); the numbers are per call, H2D / D2H included (these entry points have no device-pointer variant yet)."""
import json
import sys
import time
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from sph_b200 import capi  # noqa: E402


def timed(fn, reps=3):
    fn()
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        out = fn()
        ts.append(time.perf_counter() - t0)
    return min(ts), out


def main():
    n, k = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (512 * 217, 91)
    rng = np.random.default_rng(0)
    off = np.cumsum(rng.integers(1, max(2, n // (2 * k)), (n, k - 1)), axis=1)
    ids = np.concatenate([np.arange(n)[:, None], (np.arange(n)[:, None] + off) % n], axis=1).astype(np.int64)
    dist = np.concatenate([np.zeros((n, 1)), np.sort(rng.uniform(0.05, 4.0, (n, k - 1)), axis=1)], axis=1).astype(np.float32)
    capi.init(0)
    res = {"n": n, "k": k}
    nnz = n * k

    t, (srp, sids, sdist) = timed(lambda: capi.symmetrize_graph(ids, dist))
    res["symmetrize_graph"] = {"ms": 1e3 * t, "nnz_out": int(srp[-1]), "algorithmic_GBps": (12 * nnz + 12 * int(srp[-1])) / t / 1e9}
    t, (num, labels) = timed(lambda: capi.weak_components(sids, srp))
    res["weak_components"] = {"ms": 1e3 * t, "components": num}
    parents = (np.arange(n) // 4).astype(np.uint64)
    nm = int(parents.max()) + 1
    t, (mrp, mids, mdist) = timed(lambda: capi.merge_graph_nodes(ids, dist, parents, nm))
    res["merge_graph_nodes"] = {"ms": 1e3 * t, "nnz_out": int(mrp[-1]), "algorithmic_GBps": (12 * nnz + 12 * int(mrp[-1])) / t / 1e9}
    t, (prp, pc, pv) = timed(lambda: capi.prow_graph_ragged(mrp, mids, mdist))
    res["prow_graph_ragged"] = {"ms": 1e3 * t, "rows": nm, "max_row": int(np.diff(mrp).max())}
    cols, vals = capi.prow_gaussian_graph(ids, dist)
    rp0 = np.arange(n + 1, dtype=np.int64) * (k - 1)
    wrp, wc, wv = capi.random_walks(rp0, cols.reshape(-1), vals.reshape(-1), 90, 15, capi.W_NORMAL, seed=1, cap=n * 90 * 15 // 2)
    m1 = capi.rw_merge(wrp, wc, wv, parents, nm)
    t, _ = timed(lambda: capi.rw_as_knn_distances(*m1))
    res["rw_as_knn_distances"] = {"ms": 1e3 * t, "nnz": int(m1[0][-1])}
    t, _ = timed(lambda: capi.remove_zeros(*m1))
    res["remove_zeros"] = {"ms": 1e3 * t}
    for mode in (0, 1):
        t, _ = timed(lambda: capi.level_knn_walks(*m1, 91, mode), reps=1)
        res[f"level_knn_walks_mode{mode}"] = {"ms": 1e3 * t, "rows": nm, "k": 91}
    comp_ptr = np.arange(0, n + 1, 4, dtype=np.int64)
    if comp_ptr[-1] != n:
        comp_ptr = np.append(comp_ptr, n)
    comp_pix = np.arange(n, dtype=np.int64)
    t, _ = timed(lambda: capi.level_knn_overlap(ids, comp_ptr, comp_pix, 91), reps=1)
    res["level_knn_overlap"] = {"ms": 1e3 * t, "rows": len(comp_ptr) - 1, "k": 91}
    x = rng.standard_normal((n, 204)).astype(np.float32)
    npairs = min(200_000, 5 * n)
    pa = rng.integers(0, len(comp_ptr) - 1, npairs).astype(np.int64)
    pb = np.minimum(pa + rng.integers(1, 4, npairs), len(comp_ptr) - 2).astype(np.int64)
    t, _ = timed(lambda: capi.euclid_hausdorff(x, comp_ptr, comp_pix, pa, pb))
    res["euclid_hausdorff"] = {"ms": 1e3 * t, "pairs": npairs, "pixel_pairs": int(16 * npairs), "d": 204}
    print(json.dumps(res))


if __name__ == "__main__":
    main()
