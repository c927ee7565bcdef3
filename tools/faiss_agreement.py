"""Row-level agreement between the exact contract (fp64 pair keys, oracle orc_knn_exact_rows) and the faiss-style fp32 path
(|x|^2 + |y|^2 - 2 x.y through sgemm, oracle orc_knn_blas_l2 — the algorithm the reference's faiss call runs,
sph/utils/Knn.cpp:56-136).  For every row whose neighbour SET differs, every id in the symmetric difference must lie within
fp32 sgemm rounding of the row's k-th exact distance: |d_j - d_(k)| <= 8 * 2^-24 * (|x_i|^2 + |x_j|^2 + 2 |x_i||x_j|)
(norms + dot are each rounded to fp32 and cancel).  CPU only.  argv: workload [rows]"""
import json
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))
import bench  # noqa: E402
import oracle_lib  # noqa: E402
from sph_b200 import synth  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "C1"
nrows = int(sys.argv[2]) if len(sys.argv) > 2 else 0
x = bench.make_stack(synth, name)
n, d = x.shape
k = 91
orc = oracle_lib.load()
assert orc.blas_init()
rng = np.random.default_rng(99)
if nrows and nrows < n:
    q0 = int(rng.integers(0, n - nrows))
    rows = np.arange(q0, q0 + nrows)
else:
    q0, rows = 0, np.arange(n)
bi, bd = orc.knn_blas_l2(x, k, q0, q0 + len(rows))
ei, ed = orc.knn_exact(x, k, 1, rows=rows)
norms = (x.astype(np.float64) ** 2).sum(1)
set_diff_rows = 0
order_diff_rows = 0
explained = 0
worst_ratio = 0.0
for r in range(len(rows)):
    if np.array_equal(bi[r], ei[r]):
        continue
    sb, se = set(bi[r].tolist()), set(ei[r].tolist())
    if sb == se:
        order_diff_rows += 1
        continue
    set_diff_rows += 1
    i = rows[r]
    dk = float(ed[r, k - 1])
    ok = True
    for j in sb ^ se:
        dj = float(((x[i].astype(np.float64) - x[j].astype(np.float64)) ** 2).sum())
        tol = 8 * 2.0 ** -24 * (norms[i] + norms[j] + 2 * np.sqrt(norms[i] * norms[j]))
        worst_ratio = max(worst_ratio, abs(dj - dk) / tol)
        ok &= abs(dj - dk) <= tol
    explained += int(ok)
print(json.dumps({"workload": name, "rows": int(len(rows)), "k": k, "rows_same_order": int(len(rows) - set_diff_rows - order_diff_rows),
                  "rows_same_set_other_order": order_diff_rows, "rows_set_differs": set_diff_rows,
                  "set_differences_within_fp32_sgemm_rounding": explained, "worst_gap_over_bound": worst_ratio,
                  "max_rel_distance_diff_excluding_self": float(np.max((np.abs(np.sort(bd, 1) - ed) / np.maximum(ed, 1e-12))[:, 1:]))}))
