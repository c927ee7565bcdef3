"""Timing probe for SURVEY 8(f)-4 on the configuration that uses it (C2: 512 x 217 x 204, GEO_CENTROID): the level-0
call of ImageHierarchy::computeDistances — every pixel against its 4 spatial neighbours — through sphb_geodesic_distances, host
buffers (the 121 MB graph upload is inside the timed call).  No torch import.  Results are appended line by line to --out so
that a call cut short still leaves what it finished.

  python tools/geodesic_probe.py --out gpurun_out/r02j_geodesic_probe.jsonl [--rows 64] [--cpu-pairs 1500]"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))


def grid_pairs(h, w, r0, r1):
    p = (np.arange(r0, r1)[:, None] * w + np.arange(w)[None, :]).reshape(-1)
    r, c = np.divmod(p, w)
    pa, pb = [], []
    for dr, dc in ((-1, 0), (0, 1), (1, 0), (0, -1)):  # pixelNeighborIDs, FOUR (sph/utils/ImageHelper.hpp:13-19)
        ok = (r + dr >= 0) & (r + dr < h) & (c + dc >= 0) & (c + dc < w)
        pa.append(p[ok])
        pb.append((r[ok] + dr) * w + c[ok] + dc)
    return np.concatenate(pa).astype(np.int64), np.concatenate(pb).astype(np.int64)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="")
    ap.add_argument("--rows", type=int, default=64, help="image rows of the first, short run")
    ap.add_argument("--cpu-pairs", type=int, default=1500)
    ap.add_argument("--shape", default="512,217,204,16,1002")
    ap.add_argument("--k", type=int, default=91)
    ap.add_argument("--level1", action="store_true")
    args = ap.parse_args()
    from sph_b200 import capi, synth

    def emit(d):
        line = json.dumps(d)
        print(line, flush=True)
        if args.out:
            with open(args.out, "a") as f:
                f.write(line + "\n")

    h, w, d, M, seed = (int(v) for v in args.shape.split(","))
    t0 = time.perf_counter()
    x = synth.hyperspectral(h, w, d, endmembers=M, seed=seed).reshape(h * w, d)
    capi.init(0)
    t1 = time.perf_counter()
    ids, dist = capi.knn_exact(x, args.k)
    dist = np.sqrt(dist)  # L2squared = false: true L2 distances
    t2 = time.perf_counter()
    emit({"what": "setup", "n": h * w, "d": d, "k": args.k, "synth_s": t1 - t0, "knn_s": t2 - t1})
    for name, (r0, r1) in (("band", (0, min(args.rows, h))), ("level0", (0, h))):
        pa, pb = grid_pairs(h, w, r0, r1)
        t = time.perf_counter()
        out = capi.geodesic_distances(ids, dist, pa, pb)
        sec = time.perf_counter() - t
        st = capi.geodesic_last_stats()
        emit({"what": name, "pairs": len(pa), "sources": st["sources"], "searches": st["searches"], "expanded": st["expanded"], "passes": st["passes"],
              "seconds_host_buffers": sec, "pairs_per_s": len(pa) / sec, "no_path": int((out < 0).sum()), "checksum": float(out[out >= 0].astype(np.float64).sum())})
    if args.level1:  # level 1 of a 2 x 2 block hierarchy: Hausdorff over the blocks' pixels, every block against its 4 block neighbours
        hb, wb = h // 2, w // 2
        blk = (np.arange(h)[:, None] // 2 * wb + np.arange(w)[None, :] // 2)[: hb * 2, : wb * 2].reshape(-1)
        pix = (np.arange(hb * 2)[:, None] * w + np.arange(wb * 2)[None, :]).reshape(-1)
        order = np.argsort(blk, kind="stable")
        comp_pix = pix[order].astype(np.int64)
        comp_ptr = np.arange(hb * wb + 1, dtype=np.int64) * 4
        qa, qb = grid_pairs(hb, wb, 0, hb)
        t = time.perf_counter()
        o1 = capi.geodesic_hausdorff(ids, dist, comp_ptr, comp_pix, qa, qb)
        sec = time.perf_counter() - t
        st = capi.geodesic_last_stats()
        emit({"what": "level1_2x2_blocks", "pairs": len(qa), "components": hb * wb, "sources": st["sources"], "searches": st["searches"],
              "expanded": st["expanded"], "passes": st["passes"], "seconds_host_buffers": sec, "pairs_per_s": len(qa) / sec,
              "checksum": float(o1.astype(np.float64).sum())})
        if args.cpu_pairs > 0:
            import oracle_lib

            orc = oracle_lib.load()
            sel = np.random.default_rng(1).choice(len(qa), min(args.cpu_pairs, len(qa)), replace=False)
            t = time.perf_counter()
            ref = orc.geodesic_hausdorff(None, ids, dist, comp_ptr, comp_pix, qa[sel], qb[sel])
            csec = time.perf_counter() - t
            emit({"what": "cpu_port_sample_level1", "pairs": len(sel), "seconds": csec, "pairs_per_s": len(sel) / csec, "cores": os.cpu_count(),
                  "mismatches_vs_device": int((ref != o1[sel]).sum())})
    if args.cpu_pairs > 0:  # the oracle's restatement of computeShortestPath on a sample of the same pairs, all host cores
        import oracle_lib

        orc = oracle_lib.load()
        sel = np.random.default_rng(0).choice(len(pa), min(args.cpu_pairs, len(pa)), replace=False)
        t = time.perf_counter()
        ref = orc.shortest_paths(None, ids, dist, pa[sel], pb[sel])
        csec = time.perf_counter() - t
        emit({"what": "cpu_port_sample", "pairs": len(sel), "seconds": csec, "pairs_per_s": len(sel) / csec, "cores": os.cpu_count(),
              "mismatches_vs_device": int((ref != out[sel]).sum())})


if __name__ == "__main__":
    main()
