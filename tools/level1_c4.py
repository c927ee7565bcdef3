"""north_star target run: exact kNN graph + level-0 P rows + LEVEL-1 similarities and P matrix for the 2048 x 2048 x 128 stack
(C4; level 1 = 2 x 2 pixel blocks = 1 048 576 components) on one B200, everything resident in HBM.

The level-0 walk matrix of C4 holds ~2.2e9 entries, more than the 32-bit record counts of the sort-based passes take in one call
(DESIGN.md 8), so RW1 (doRandomWalks) and RW2 (mergeNodesRandomWalks) run per BAND of image rows: with 2 x 2 blocks a parent never
straddles a band of an even number of rows, so the merged rows of a band depend on that band's walks only and the bands' merged
rows concatenate to exactly the one-call result.  D1 -> B3 -> B7 -> B6 (sphb_level_prob_dist_dev) then run once over all
1 048 576 merged rows (SLAB accumulator flavour of sims.cu).

Parity: --parity-rows sampled rows of createSimilarities (D1) + in-place Gaussian rows (B3) recomputed by the oracle from the
SAME merged matrix (orc_create_similarities_rows on a row range), columns bit-exact, values rtol 2e-5.

  python tools/level1_c4.py --out gpurun_out/r02_level1_c4.json            (C4, ~1 min of GPU time + the oracle sample)
  python tools/level1_c4.py --workload C2 --band-rows 64                   (small: checks banded == one call)
"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="C4")
    ap.add_argument("--band-rows", type=int, default=64, help="image rows per band (even)")
    ap.add_argument("--parity-rows", type=int, default=256)
    ap.add_argument("--out", default="")
    ap.add_argument("--shape", default="", help="h,w,d,endmembers,seed instead of a named workload")
    ap.add_argument("--skip-knn-timing", action="store_true")
    args = ap.parse_args()

    import torch
    from sph_b200 import capi, host, synth
    import bench

    if args.shape:
        bench.SHAPES["custom"] = tuple(int(v) for v in args.shape.split(","))
        args.workload = "custom"
    h, w, d, M, seed = bench.SHAPES[args.workload]
    k = bench.K_NEIGHBOURS
    n = h * w
    assert args.band_rows % 2 == 0
    capi.init(0)
    torch.cuda.set_device(0)
    s = torch.cuda.current_stream().cuda_stream

    def ev():
        return torch.cuda.Event(enable_timing=True)

    def timed(fn):
        e0, e1 = ev(), ev()
        e0.record()
        r = fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1), r

    # ---- Stage A + B2, resident
    x = torch.from_numpy(bench.make_stack(synth, args.workload)).cuda()
    ids = torch.empty((n, k), dtype=torch.int64, device="cuda")
    dist = torch.empty((n, k), dtype=torch.float32, device="cuda")
    cols = torch.empty((n, k - 1), dtype=torch.int32, device="cuda")
    vals = torch.empty((n, k - 1), dtype=torch.float32, device="cuda")
    knn_ms, _ = timed(lambda: capi.knn_exact_dev(x.data_ptr(), n, d, k, capi.METRIC_L2, False, 0, n, ids.data_ptr(), dist.data_ptr(), s))
    prow_ms, _ = timed(lambda: capi.prow_gaussian_graph_dev(ids.data_ptr(), dist.data_ptr(), n, k, 0.0, cols.data_ptr(), vals.data_ptr(), s))
    del x, ids, dist
    torch.cuda.empty_cache()

    # ---- level 0 -> level 1 hierarchy: 2 x 2 pixel blocks
    W, L = 90, 15
    rp0 = torch.arange(n + 1, dtype=torch.int64, device="cuda") * (k - 1)
    dc0, dv0 = cols.reshape(-1), vals.reshape(-1)
    rows, colsi = np.divmod(np.arange(n), w)
    hb, wb = (h + 1) // 2, (w + 1) // 2
    parents = ((rows // 2) * wb + colsi // 2).astype(np.int64)
    n1 = hb * wb
    dpar = torch.from_numpy(parents).cuda()
    _, ks = host.level_schedule([n, n1], k)
    k1 = int(ks[1])

    band = args.band_rows * w
    cap_w = band * W * L  # a start point visits at most W * L distinct nodes
    wrp_b = torch.empty(band + 1, dtype=torch.int64, device="cuda")
    wc = torch.empty(cap_w, dtype=torch.int32, device="cuda")
    wv = torch.empty(cap_w, dtype=torch.float32, device="cuda")
    wrp_full = torch.empty(n + 1, dtype=torch.int64, device="cuda")
    mrp_b = torch.empty(n1 + 1, dtype=torch.int64, device="cuda")
    mc_b = torch.empty(cap_w, dtype=torch.int32, device="cuda")
    mv_b = torch.empty(cap_w, dtype=torch.float32, device="cuda")
    est = int(n1 * 4 * 1350 // 2)  # a merged row holds at most 4 x W x L entries; half of that is ample
    mc = torch.empty(est, dtype=torch.int32, device="cuda")
    mv = torch.empty(est, dtype=torch.float32, device="cuda")
    mrp = torch.zeros(n1 + 1, dtype=torch.int64, device="cuda")
    rw_ms = mg_ms = 0.0
    walk_nnz = merged = 0
    for q0 in range(0, n, band):
        q1 = min(n, q0 + band)
        ms, wn = timed(lambda: capi.random_walks_dev(n, rp0.data_ptr(), dc0.data_ptr(), dv0.data_ptr(), W, L, capi.W_NORMAL, 1, q0, q1, wrp_b.data_ptr(),
                                                     wc.data_ptr(), wv.data_ptr(), cap_w, s))
        rw_ms += ms
        walk_nnz += wn
        # the band's walks as a full-height CSR (rows outside the band empty): row and column ids stay global
        wrp_full[: q0 + 1] = 0
        wrp_full[q0: q1 + 1] = wrp_b[: q1 - q0 + 1]
        wrp_full[q1 + 1:] = wn
        ms, mn = timed(lambda: capi.rw_merge_dev(n, wrp_full.data_ptr(), wc.data_ptr(), wv.data_ptr(), wn, dpar.data_ptr(), n1, True, True,
                                                 mrp_b.data_ptr(), mc_b.data_ptr(), mv_b.data_ptr(), cap_w, s))
        mg_ms += ms
        p0, p1 = (q0 // w // 2) * wb, ((q1 // w + 1) // 2) * wb
        assert int(mrp_b[p0]) == 0 and int(mrp_b[p1]) == mn, "a band's merged rows must lie inside the band's parents"
        assert merged + mn <= est
        mc[merged: merged + mn] = mc_b[:mn]
        mv[merged: merged + mn] = mv_b[:mn]
        mrp[p0 + 1: p1 + 1] = mrp_b[p0 + 1: p1 + 1] + merged
        merged += mn
    del wc, wv, mc_b, mv_b, wrp_full, wrp_b, mrp_b
    torch.cuda.empty_cache()

    # column lengths of the merged matrix -> product count of the Gustavson expansion (D1's work)
    lens = torch.zeros(n1, dtype=torch.int64, device="cuda")
    for c0 in range(0, merged, 1 << 28):
        lens += torch.bincount(mc[c0: min(merged, c0 + (1 << 28))].long(), minlength=n1)
    products = float((lens.double() ** 2).sum())
    del lens
    out = {"workload": f"{args.workload}: {h}x{w}x{d}, k = {k}; level 1 = 2x2 pixel blocks ({n1} components), k_1 = {k1}, walks W={W} L={L} NORMAL, "
                       f"RW1/RW2 in bands of {args.band_rows} image rows, resident in HBM",
           "knn_ms": knn_ms, "prow_level0_ms": prow_ms, "random_walks_level0_ms": rw_ms, "rw_merge_ms": mg_ms,
           "walk_nnz": int(walk_nnz), "merged_nnz": int(merged), "merged_entries_per_component": merged / n1, "d1_products": products,
           "walk_steps": n * W * L, "walk_steps_per_s": n * W * L / (rw_ms * 1e-3)}
    if merged >= 2 ** 31:
        out["level1_prob_dist_ms"] = None
        out["note"] = ("the merged walk matrix holds more than 2^31 - 1 entries: createSimilarities (D1) counts records in 32 bits and is not run; "
                       f"its Gustavson expansion would be {products:.3g} products")
        line = json.dumps(out)
        print(line)
        if args.out:
            with open(args.out, "w") as f:
                f.write(line + "\n")
        return
    # ---- the level: D1 -> CSR -> B3 -> B7 -> B6 in one call
    cap_p = 2 * n1 * k1
    prp = torch.empty(n1 + 1, dtype=torch.int64, device="cuda")
    pc = torch.empty(cap_p, dtype=torch.int32, device="cuda")
    pv = torch.empty(cap_p, dtype=torch.float32, device="cuda")
    lv_ms, pnnz = timed(lambda: capi.level_prob_dist_dev(n1, mrp.data_ptr(), mc.data_ptr(), mv.data_ptr(), merged, k1, 1e-4, 0, None, True, 0.0, 1, True,
                                                         prp.data_ptr(), pc.data_ptr(), pv.data_ptr(), cap_p, s))
    stage = dict(capi.level_last_timing())
    out.update({"level1_prob_dist_ms": lv_ms, "level1_stages_ms": stage, "total_ms": knn_ms + prow_ms + rw_ms + mg_ms + lv_ms, "p_nnz": int(pnnz)})

    # ---- parity of D1 + B3 on sampled rows against the oracle (same merged matrix)
    if args.parity_rows > 0:
        orc = bench.load_oracle(os.cpu_count())
        sc = torch.empty((n1, k1), dtype=torch.int32, device="cuda")
        sv = torch.empty((n1, k1), dtype=torch.float32, device="cuda")
        snz = torch.empty(n1, dtype=torch.int32, device="cuda")
        d1_ms, _ = timed(lambda: capi.create_similarities_dev(n1, mrp.data_ptr(), mc.data_ptr(), mv.data_ptr(), merged, k1, 1e-4, 0, None, sc.data_ptr(),
                                                              sv.data_ptr(), snz.data_ptr(), s))
        out["create_similarities_alone_ms"] = d1_ms
        h_mrp, h_mc, h_mv = mrp.cpu().numpy(), mc[:merged].cpu().numpy().view(np.uint32), mv[:merged].cpu().numpy()
        r0 = (n1 // 2 // wb) * wb + wb // 3
        r1 = r0 + args.parity_rows
        t0 = time.perf_counter()
        oc, ov, onz = orc.create_similarities_rows(h_mrp, h_mc, h_mv, k1, r0, r1, 1e-4, 0)
        osec = time.perf_counter() - t0
        gc, gv, gnz = sc[r0:r1].cpu().numpy().view(np.uint32), sv[r0:r1].cpu().numpy(), snz[r0:r1].cpu().numpy()
        same = 0
        worst = 0.0
        for i in range(r1 - r0):
            m = int(onz[i])
            if gnz[i] == m and np.array_equal(gc[i, :m], oc[i, :m]):
                same += 1
                if m:
                    worst = max(worst, float((np.abs(gv[i, :m] - ov[i, :m]) / np.maximum(np.abs(ov[i, :m]), 1e-30)).max()))
        out["parity"] = {"rows_checked": int(r1 - r0), "first_row": int(r0), "rows_identical_pattern": int(same), "max_rel_on_identical_rows": worst,
                         "oracle_seconds": osec, "what": "createSimilarities rows (columns, -ln values normalised) vs orc_create_similarities_rows on the same merged walks"}
    line = json.dumps(out)
    print(line)
    if args.out:
        with open(args.out, "w") as f:
            f.write(line + "\n")


if __name__ == "__main__":
    main()
