"""Level-1 pipeline probe at a C2-shaped size through the host-buffer C ABI: one call of every graph pass
(walks, merge, similarities, in-place Gaussian rows, symmetrise) so that an ncu launch list of this command
gives per-kernel time and DRAM bytes.  argv: [h w d k] (default C2: 512 217 204 91), env PROBE_REPS."""
import json
import os
import sys
import time
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from sph_b200 import capi, host, synth  # noqa: E402


def main():
    h, w, d, k = (int(a) for a in sys.argv[1:5]) if len(sys.argv) > 4 else (512, 217, 204, 91)
    reps = int(os.environ.get("PROBE_REPS", "1"))
    capi.init(0)
    x = synth.hyperspectral(h, w, d, endmembers=16, seed=1002)
    n = h * w
    ids, dist, cols, vals = capi.knn_graph_prow(x, k, 0.0)
    rp0 = np.arange(n + 1, dtype=np.int64) * (k - 1)
    rows, colsi = np.divmod(np.arange(n), w)
    hb, wb = (h + 1) // 2, (w + 1) // 2
    parents = ((rows // 2) * wb + colsi // 2).astype(np.uint64)
    n1 = hb * wb
    _, ks = host.level_schedule([n, n1], k)
    k1 = ks[1]
    res = {"n": n, "n1": n1, "k1": int(k1)}
    for r in range(reps):
        t = {}
        t0 = time.perf_counter()
        wrp, wc, wv = capi.random_walks(rp0, cols.reshape(-1), vals.reshape(-1), 90, 15, capi.W_NORMAL, seed=1, cap=n * 90 * 15 // 2)
        t["random_walks"] = time.perf_counter() - t0
        t0 = time.perf_counter()
        mrp, mc, mv = capi.rw_merge(wrp, wc, wv, parents, n1, norm=True, weight_by_size=True)
        t["rw_merge"] = time.perf_counter() - t0
        t0 = time.perf_counter()
        sc, sv, snz = capi.create_similarities(mrp, mc, mv, k1, 1e-4)
        t["create_similarities"] = time.perf_counter() - t0
        prp = np.concatenate([[0], np.cumsum(snz, dtype=np.int64)])
        keep = np.arange(sc.shape[1])[None, :] < snz[:, None]
        pc, pv = sc[keep], sv[keep]
        t0 = time.perf_counter()
        pv = capi.prow_gaussian_inplace(prp, pv, 0.0, -1)
        t["prow_inplace"] = time.perf_counter() - t0
        t0 = time.perf_counter()
        orp, oc, ov = capi.symmetrize(prp, pc, pv)
        t["symmetrize"] = time.perf_counter() - t0
        t0 = time.perf_counter()
        frp, fc, fv = capi.level_prob_dist(mrp, mc, mv, int(k1), 1e-4, 0, None, True, 0.0, 1, True)
        t["fused_level_prob_dist_host"] = time.perf_counter() - t0
        res[f"fused_stage_ms_rep{r}"] = capi.level_last_timing()
        res[f"fused_equals_staged_rep{r}"] = bool(np.array_equal(frp, orp) and np.array_equal(fc, oc) and np.array_equal(fv, ov))
        res[f"wall_ms_rep{r}"] = {a: 1e3 * b for a, b in t.items()}
    keepm = mv > 1e-4 * 1e-5
    lens = np.bincount(mc[keepm].astype(np.int64), minlength=n1).astype(np.float64)
    res["sims_products"] = float((lens ** 2).sum())
    res["sims_col_len_max"] = float(lens.max())
    res["sims_col_len_mean"] = float(lens.mean())
    res.update(walk_nnz=int(wrp[-1]), merged_nnz=int(mrp[-1]), sim_nnz=int(prp[-1]), p_nnz=int(orp[-1]))
    print(json.dumps(res))


if __name__ == "__main__":
    main()
