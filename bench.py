#!/usr/bin/env python
"""bench.py — headline benchmark of the SPH hot path on B200.

A "step" is one pass of the hot path over one synthetic image stack:
exact kNN graph (Stage A, tcgen05 sweep + exact re-rank) followed by the level-0
perplexity rows (Stage B) for every pixel; query rows are sharded over the ranks
(database replicated, no collective inside the timed resident step), total work is
fixed, so `scaling` is "strong".

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload C4] [--impl reference]

Prints ONE JSON line (rank 0).  See DESIGN.md §5 for every field.
"""
from __future__ import annotations

import argparse
import contextlib
import ctypes
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

METRIC = "knn_graph_plus_similarity_pixels_per_s"
UNIT = "pixels/s"
K_NEIGHBOURS = 91  # nKnns = 90 + self (evaluation/RunEvaluation.cpp:241-248, comparison/*.json "nKnns")

SHAPES = {  # name: (h, w, d, endmembers, seed)   
    "C1": (145, 145, 200, 16, 1001),
    "C2": (512, 217, 204, 16, 1002),
    "C3": (1080, 1920, 3, 0, 1003),
    "C4": (2048, 2048, 128, 32, 1004),
    "C5": (1024, 1024, 1000, 64, 1005),
    "tiny": (128, 128, 128, 8, 7),
}


def make_stack(synth, name):
    h, w, d, M, seed = SHAPES[name]
    if d == 3:
        return synth.rgb(h, w, seed=seed)
    return synth.hyperspectral(h, w, d, endmembers=M, seed=seed)


def _peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        j = json.loads(p.read_text())
        return dict(hbm=j.get("hbm_gbs", 6650.0), tflops=j.get("bf16_tflops_sustained", 1400.0), tflops_burst=j.get("bf16_tflops", 1590.0), src="measured")
    return dict(hbm=6650.0, tflops=1400.0, tflops_burst=1590.0, src="fallback")


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons during the timed region."""

    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index: int):
        self.index = index
        self.samples = []
        self._stop = threading.Event()
        self._t = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i", str(self.index)],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.samples.append([s.strip() for s in out.split(",")])
            except Exception:
                pass
            self._stop.wait(0.2)

    def __enter__(self):
        self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        self._t.join(timeout=10)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        sm = sorted(float(s[0]) for s in self.samples if s[0].replace(".", "").isdigit())
        mx = max(float(s[1]) for s in self.samples if s[1].replace(".", "").isdigit())
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for s in self.samples:
            for nm, v in zip(names, s[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(self.samples)}


def load_oracle(threads=None):
    """The CPU checker / baseline (oracle/): only the parity block, the cpu_baseline legs and --impl reference use it.
    threads: OpenMP team size of this process (torchrun exports OMP_NUM_THREADS=1; the ranks share the host cores)."""
    sys.path.insert(0, str(ROOT / "tests"))
    import oracle_lib

    orc = oracle_lib.load()
    if threads:
        try:
            ctypes.CDLL("libgomp.so.1").omp_set_num_threads(int(threads))
        except OSError:
            pass
    return orc


def cpu_reference_step(orc, x, k, q0, q1):
    """The reference's CPU path for a slab of query rows: faiss-style blocked sgemm + heaps
    (oracle/oracle_knn.cpp orc_knn_blas_l2), post-passes, level-0 Gaussian rows."""
    ids, dist = orc.knn_blas_l2(x, k, q0, q1)
    orc.knn_postprocess(ids, dist, False, np.arange(q0, q1))
    orc.prow_gaussian_graph(ids, dist, 0.0)
    return ids, dist


def cpu_sgemm_info(orc, d, steps_stats):
    """sgemm rate achieved inside the arm (mean over its threads) next to a standalone sgemm of the same block shape."""
    flop = sum(s["sgemm_flop"] for s in steps_stats)
    secs = sum(s["sgemm_s"] for s in steps_stats)
    heap = sum(s["heap_s"] for s in steps_stats)
    out = {"sgemm_gflops": flop / secs / 1e9 if secs > 0 else None, "sgemm_seconds": secs, "heap_seconds": heap,
           "threads": steps_stats[-1]["threads"] if steps_stats else None}
    try:
        out["sgemm_standalone_gflops"] = orc.sgemm_standalone_gflops(d, 1.0)
    except Exception:
        out["sgemm_standalone_gflops"] = None
    return out


def run_cpu_sample(orc, x, k, budget_s, slab):
    if not orc.blas_init():
        raise RuntimeError("OpenBLAS (numpy.libs) not found for the CPU baseline")
    n = x.shape[0]
    done = 0
    stats = []
    t0 = time.perf_counter()
    q = 0
    while True:
        q1 = min(n, q + slab)
        cpu_reference_step(orc, x, k, q, q1)
        stats.append(orc.knn_blas_last_stats())
        done += q1 - q
        q = q1 % n
        el = time.perf_counter() - t0
        if el >= budget_s or done >= n:
            break
    return done, el, stats


@contextlib.contextmanager
def stdout_to_stderr():
    """Route file descriptor 1 (Python's and the C library's stdout) to stderr inside the block."""
    sys.stdout.flush()
    saved = os.dup(1)
    os.dup2(2, 1)
    try:
        yield
    finally:
        try:
            sys.stdout.flush()
            ctypes.CDLL(None).fflush(None)
        finally:
            os.dup2(saved, 1)
            os.close(saved)


# ---------------------------------------------------------------------------------------------------------------------
# parity of the benchmarked configuration (checked AFTER the timed region, never inside it)
# ---------------------------------------------------------------------------------------------------------------------
def parity_rows(capi, n, q0, q1, target, seed):
    """Global ids of the rows to check: a seeded random sample of this rank's slab plus rows of the last kNN call that took
    the overflow re-sweep (3-product filter) and the exact SIMT fallback."""
    rng = np.random.default_rng(seed)
    nq = q1 - q0
    rows = q0 + rng.choice(nq, size=min(target, nq), replace=False)
    resweep = capi.knn_last_overflow_rows(0)
    simt = capi.knn_last_overflow_rows(1)
    extra_r = resweep[rng.choice(len(resweep), size=min(32, len(resweep)), replace=False)] if len(resweep) else np.zeros(0, np.int64)
    extra_s = simt[rng.choice(len(simt), size=min(32, len(simt)), replace=False)] if len(simt) else np.zeros(0, np.int64)
    allrows = np.unique(np.concatenate([rows, extra_r, extra_s]).astype(np.int64))
    return allrows, int(len(extra_r)), int(len(extra_s)), int(len(resweep)), int(len(simt))


def parity_check(orc, x, k, rows, got):
    """got = (ids, dist, cols, vals) of `rows` from the CUDA path; the oracle recomputes them from scratch:
    orc_knn_exact_rows + orc_knn_postprocess_rows (oracle/oracle_knn.cpp) and orc_prow_gaussian_graph (oracle/oracle_prow.cpp)."""
    ids, dist, cols, vals = got
    t0 = time.perf_counter()
    rids, rdist = orc.knn_reference_graph(x, k, rows=rows)
    rcols, rvals, _ = orc.prow_gaussian_graph(rids, rdist, 0.0)
    id_bad = int((ids != rids).any(axis=1).sum())
    dist_bad = int((dist != rdist).sum())
    col_bad = int((cols != rcols).any(axis=1).sum())
    ok = (cols == rcols).all(axis=1)
    denom = np.maximum(np.abs(rvals[ok]), 1e-30)
    rel = float((np.abs(vals[ok] - rvals[ok]) / denom).max()) if ok.any() else 0.0
    return {"rows_checked": int(len(rows)), "id_mismatch_rows": id_bad, "dist_mismatch": dist_bad, "prow_col_mismatch_rows": col_bad,
            "prow_max_rel": rel, "oracle_seconds": time.perf_counter() - t0}


# ---------------------------------------------------------------------------------------------------------------------
# second half of BASELINE.json's metric: level-1 P matrix on a FIXED hierarchy, resident in HBM
# ---------------------------------------------------------------------------------------------------------------------
def level1_p_matrix(capi, synth, orc, peaks, workload="C2", cpu=True):
    """Level 0 (kNN + P rows) is input.  Level 1 = 2x2 pixel blocks (the ImageHierarchy merge loop is host code and out of
    scope).  Timed with CUDA events on device-resident matrices (the `_dev` entry points):
      RW1 doRandomWalks W=90 L=15 NORMAL (sph/utils/SparseMatrixAlgorithms.cpp:34-290)
      RW2 mergeNodesRandomWalks (:292-352)
      D1 createSimilarities(k_1, 1e-4) -> B3 in-place Gaussian rows (ignore = -1) -> B7 zero removal -> B6 symmetrizeTSNE
         = sphb_level_prob_dist_dev (sph/LevelSimilarities.cpp:460-508, 566-581, 589-623)
    and, beside it, the oracle's same stages on the same inputs on the host cores (cpu_baseline) with a parity verdict."""
    import torch
    from sph_b200 import host

    h, w, d, M, seed = SHAPES[workload]
    k = K_NEIGHBOURS
    x = make_stack(synth, workload)
    n = h * w
    ids, dist, cols, vals = capi.knn_graph_prow(x, k, 0.0)
    del x
    s = torch.cuda.current_stream().cuda_stream
    W, L = 90, 15
    rp0 = torch.arange(n + 1, dtype=torch.int64, device="cuda") * (k - 1)
    dc0 = torch.from_numpy(cols.reshape(-1).view(np.int32)).cuda()
    dv0 = torch.from_numpy(vals.reshape(-1)).cuda()
    cap_w = n * W * L // 2
    wrp = torch.empty(n + 1, dtype=torch.int64, device="cuda")
    wc = torch.empty(cap_w, dtype=torch.int32, device="cuda")
    wv = torch.empty(cap_w, dtype=torch.float32, device="cuda")
    rows, colsi = np.divmod(np.arange(n), w)
    hb, wb = (h + 1) // 2, (w + 1) // 2
    parents = ((rows // 2) * wb + colsi // 2).astype(np.uint64)
    n1 = hb * wb
    dpar = torch.from_numpy(parents.view(np.int64)).cuda()
    _, ks = host.level_schedule([n, n1], k)
    k1 = int(ks[1])

    def timed(fn, reps=3):
        out = fn()  # warm-up (workspace cache)
        best = None
        for _ in range(reps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            out = fn()
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1)
            best = ms if best is None else min(best, ms)
        return best, out

    rw_ms, wnnz = timed(lambda: capi.random_walks_dev(n, rp0.data_ptr(), dc0.data_ptr(), dv0.data_ptr(), W, L, capi.W_NORMAL, 1, 0, n, wrp.data_ptr(),
                                                      wc.data_ptr(), wv.data_ptr(), cap_w, s), reps=2)
    mrp = torch.empty(n1 + 1, dtype=torch.int64, device="cuda")
    mc = torch.empty(wnnz, dtype=torch.int32, device="cuda")
    mv = torch.empty(wnnz, dtype=torch.float32, device="cuda")
    mg_ms, mnnz = timed(lambda: capi.rw_merge_dev(n, wrp.data_ptr(), wc.data_ptr(), wv.data_ptr(), wnnz, dpar.data_ptr(), n1, True, True, mrp.data_ptr(),
                                                  mc.data_ptr(), mv.data_ptr(), wnnz, s))
    cap_p = 2 * n1 * k1
    prp = torch.empty(n1 + 1, dtype=torch.int64, device="cuda")
    pc = torch.empty(cap_p, dtype=torch.int32, device="cuda")
    pv = torch.empty(cap_p, dtype=torch.float32, device="cuda")
    stage = {}

    def level():
        m = capi.level_prob_dist_dev(n1, mrp.data_ptr(), mc.data_ptr(), mv.data_ptr(), mnnz, k1, 1e-4, 0, None, True, 0.0, 1, True, prp.data_ptr(),
                                     pc.data_ptr(), pv.data_ptr(), cap_p, s)
        stage.update(capi.level_last_timing())
        return m

    lv_ms, pnnz = timed(level)
    # merged matrix on the host: product count of the Gustavson expansion (D1's algorithmic work) and the CPU baseline's input
    h_mrp, h_mc, h_mv = mrp.cpu().numpy(), mc[:mnnz].cpu().numpy().view(np.uint32), mv[:mnnz].cpu().numpy()
    kept = ~(np.abs(h_mv) <= np.float32(1e-4) * np.float32(1e-5))
    lens = np.bincount(h_mc[kept].astype(np.int64), minlength=n1).astype(np.float64)
    products = float((lens ** 2).sum())
    sim_nnz = None
    g_prp, g_pc, g_pv = prp.cpu().numpy(), pc[:pnnz].cpu().numpy().view(np.uint32), pv[:pnnz].cpu().numpy()

    def gbs(nbytes, ms):
        return nbytes / (ms * 1e-3) / 1e9

    def entry(name, kernel, ms, nbytes, note=None):
        e = {"pass": name, "kernel": kernel, "bound": "hbm", "kernel_ms": ms, "algorithmic_bytes_per_launch": nbytes, "achieved": gbs(nbytes, ms),
             "peak": peaks["hbm"], "unit": "GB/s", "frac": gbs(nbytes, ms) / peaks["hbm"]}
        if note:
            e["note"] = note
        return e

    passes = [
        entry("RW1 doRandomWalks", "random_walk_search_kernel (binary search in fp64 running sums, radix grouping) + compaction", rw_ms, n * W * L * 8.0 * (k - 1) / 2 + 8.0 * wnnz,
              "360 B per step (half a 90-entry row of (col, val) pairs) + 8 B per output entry, SURVEY §8d"),
        entry("RW2 mergeNodesRandomWalks", "merge_expand + radix sort + merge_reduce", mg_ms, 8.0 * wnnz + 8.0 * mnnz + 8.0 * n),
        entry("D1 createSimilarities", "sims_class_kernel (class-partitioned shared-memory accumulators; SLAB flavour above ~52 k components) + transpose sorts", stage["similarities_ms"], 8.0 * products + 8.0 * n1 * k1,
              "8 B per expanded product (row id + value of the transposed matrix) + 8 B per output slot"),
        entry("B3 Gaussian rows in place", "prow_inplace_kernel", stage["gaussian_ms"], 8.0 * n1 * k1 + 8.0 * n1),
        entry("B6 symmetrizeTSNE", "sym_expand + radix sort + sym_combine", stage["symmetrize_ms"], 8.0 * n1 * k1 + 8.0 * pnnz),
    ]
    out = {"workload": f"{workload}-shaped {h}x{w}x{d}: level 1 = 2x2 pixel blocks ({n1} components), k_1 = {k1}, walks W={W} L={L} NORMAL, resident in HBM",
           "wall_ms": lv_ms, "stages_ms": dict(stage), "rw_merge_ms": mg_ms, "random_walks_level0_ms": rw_ms,
           "walk_nnz": int(wnnz), "merged_nnz": int(mnnz), "p_nnz": int(pnnz), "d1_products": products}
    # the same level through the host-buffer C ABI (H2D of the merged matrix + D2H of P inside the wall time)
    t0 = time.perf_counter()
    hp = capi.level_prob_dist(h_mrp, h_mc, h_mv, k1, 1e-4, 0, None, True, 0.0, 1, True)
    out["wall_ms_host_buffers"] = 1e3 * (time.perf_counter() - t0)
    out["host_equals_resident"] = bool(np.array_equal(hp[0], g_prp) and np.array_equal(hp[1], g_pc) and np.array_equal(hp[2], g_pv))
    # SURVEY 8(e), D1 sharded by output row: the one-rank replay of multigpu.ShardedLevel (slab call + level tail) against the fused call
    try:
        from sph_b200.multigpu import ShardedLevel

        job = ShardedLevel(0, 1)
        sh_ms, (srp, scol, sval) = timed(lambda: job(mrp, mc[:mnnz], mv[:mnnz], k1, 0.0))
        out["sharded_level_one_rank"] = {"ms": sh_ms, "equals_fused_level": bool(torch.equal(srp, prp) and scol.numel() == pnnz and
                                                                                   torch.equal(scol, pc[:pnnz]) and torch.equal(sval, pv[:pnnz]))}
        del srp, scol, sval
    except Exception as e:  # never let a side measurement take the bench line down
        out["sharded_level_one_rank"] = {"error": repr(e)[:300]}
    if cpu and orc is not None:
        t0 = time.perf_counter()
        sc, sv, snz = orc.create_similarities(h_mrp, h_mc, h_mv, k1, 1e-4, 0)
        t1 = time.perf_counter()
        crp = np.concatenate([[0], np.cumsum(snz, dtype=np.int64)])
        keep = np.arange(sc.shape[1])[None, :] < snz[:, None]
        cc, cv = sc[keep], sv[keep]
        cv = orc.prow_gaussian_inplace(crp, cv, 0.0, -1)
        t2 = time.perf_counter()
        zrp, zc, zv = orc.remove_zeros(crp, cc, cv)
        orp, oc, ov = orc.symmetrize(zrp, zc, zv)
        t3 = time.perf_counter()
        same = 0
        worst = 0.0
        for i in range(n1):
            a, b = slice(g_prp[i], g_prp[i + 1]), slice(orp[i], orp[i + 1])
            if (g_prp[i + 1] - g_prp[i]) == (orp[i + 1] - orp[i]) and np.array_equal(g_pc[a], oc[b]):
                same += 1
                if g_prp[i + 1] > g_prp[i]:
                    worst = max(worst, float((np.abs(g_pv[a] - ov[b]) / np.maximum(np.abs(ov[b]), 1e-30)).max()))
        out["cpu_baseline"] = {"wall_ms": 1e3 * (t3 - t0), "stages_ms": {"similarities_ms": 1e3 * (t1 - t0), "gaussian_ms": 1e3 * (t2 - t1),
                                                                         "remove_zeros_symmetrize_ms": 1e3 * (t3 - t2)},
                               "cores": os.cpu_count(), "kind": "port",
                               "sample": "the whole level (all rows): oracle createSimilarities (row-wise SpGEMM, OpenMP) + in-place Gaussian rows + zero removal + symmetrizeTSNE on the same merged walks"}
        out["parity"] = {"rows": int(n1), "rows_identical_pattern": int(same), "max_rel_on_identical_rows": worst, "p_nnz_oracle": int(orp[-1]),
                         "note": "a 1-ulp tie at the k-th slot of createSimilarities changes that row's pattern (tolerated by the reference's own test, tests/UtilsTests.cpp:984-1027)"}
    return out, passes


def geodesic_block(capi, synth, orc, workload="C2", cpu_pairs=20000):
    """SURVEY 8(f)-4 on the configuration that uses it (C2, GEO_CENTROID): the geodesic component distances of level 0 (every pixel
    against its 4 spatial neighbours, ImageHierarchy::computeDistances) and of level 1 of a 2 x 2 block hierarchy, through the
    host-buffer C ABI (the graph upload is inside the wall time), with a sample of the pairs recomputed by the oracle."""
    h, w, d, M, seed = SHAPES[workload]
    k = K_NEIGHBOURS
    x = make_stack(synth, workload)
    ids, dist = capi.knn_exact(x, k)
    del x
    dist = np.sqrt(dist)  # L2squared = false: true L2 edge weights

    def grid_pairs(hh, ww):
        p = np.arange(hh * ww)
        r, c = np.divmod(p, ww)
        pa, pb = [], []
        for dr, dc in ((-1, 0), (0, 1), (1, 0), (0, -1)):  # pixelNeighborIDs, FOUR
            ok = (r + dr >= 0) & (r + dr < hh) & (c + dc >= 0) & (c + dc < ww)
            pa.append(p[ok])
            pb.append((r[ok] + dr) * ww + c[ok] + dc)
        return np.concatenate(pa).astype(np.int64), np.concatenate(pb).astype(np.int64)

    out = {"workload": f"{workload} {h}x{w}x{d}, k = {k}, true L2 edge weights; host buffers"}
    pa, pb = grid_pairs(h, w)
    capi.geodesic_distances(ids[:64], dist[:64], [0], [1])  # context / workspace warm-up on a tiny graph
    t0 = time.perf_counter()
    g0 = capi.geodesic_distances(ids, dist, pa, pb)
    sec = time.perf_counter() - t0
    st = capi.geodesic_last_stats()
    out["level0"] = {"pairs": int(len(pa)), "seconds": sec, "pairs_per_s": len(pa) / sec, "searches": st["searches"], "sources": st["sources"],
                     "vertices_expanded": st["expanded"], "passes": st["passes"], "no_path": int((g0 < 0).sum())}
    hb, wb = h // 2, w // 2
    blk = (np.arange(h)[:, None] // 2 * wb + np.arange(w)[None, :] // 2)[: hb * 2, : wb * 2].reshape(-1)
    pix = (np.arange(hb * 2)[:, None] * w + np.arange(wb * 2)[None, :]).reshape(-1)
    comp_pix = pix[np.argsort(blk, kind="stable")].astype(np.int64)
    comp_ptr = np.arange(hb * wb + 1, dtype=np.int64) * 4
    qa, qb = grid_pairs(hb, wb)
    t0 = time.perf_counter()
    g1 = capi.geodesic_hausdorff(ids, dist, comp_ptr, comp_pix, qa, qb)
    sec = time.perf_counter() - t0
    st = capi.geodesic_last_stats()
    out["level1_2x2_blocks"] = {"pairs": int(len(qa)), "components": int(hb * wb), "seconds": sec, "pairs_per_s": len(qa) / sec, "searches": st["searches"],
                                "vertices_expanded": st["expanded"], "passes": st["passes"]}
    # the same two calls on a graph kept on the device (sphb_geodesic_graph_create: one upload for all the levels of a hierarchy)
    t0 = time.perf_counter()
    with capi.GeodesicGraph(ids, dist) as gg:
        up = time.perf_counter() - t0
        t0 = time.perf_counter()
        h0 = gg.distances(pa, pb)
        r0s = time.perf_counter() - t0
        t0 = time.perf_counter()
        h1 = gg.hausdorff(comp_ptr, comp_pix, qa, qb)
        r1s = time.perf_counter() - t0
    out["resident_graph"] = {"upload_seconds": up, "level0_seconds": r0s, "level0_pairs_per_s": len(pa) / r0s, "level1_seconds": r1s,
                             "level1_pairs_per_s": len(qa) / r1s, "equals_one_shot_calls": bool(np.array_equal(h0, g0) and np.array_equal(h1, g1))}
    if orc is not None and cpu_pairs > 0:
        rng = np.random.default_rng(0)
        s0 = rng.choice(len(pa), min(cpu_pairs, len(pa)), replace=False)
        t0 = time.perf_counter()
        r0 = orc.shortest_paths(None, ids, dist, pa[s0], pb[s0])
        c0 = time.perf_counter() - t0
        s1 = rng.choice(len(qa), min(cpu_pairs, len(qa)), replace=False)
        t0 = time.perf_counter()
        r1 = orc.geodesic_hausdorff(None, ids, dist, comp_ptr, comp_pix, qa[s1], qb[s1])
        c1 = time.perf_counter() - t0
        out["cpu_baseline"] = {"kind": "port", "cores": os.cpu_count(), "level0_pairs_per_s": len(s0) / c0, "level1_pairs_per_s": len(s1) / c1,
                               "sample": f"{len(s0)} level-0 pairs and {len(s1)} level-1 pairs of the same calls (oracle/oracle_geodesic.cpp: Dijkstra per source, OpenMP)"}
        out["parity"] = {"level0_pairs_checked": int(len(s0)), "level0_mismatches": int((r0 != g0[s0]).sum()), "level1_pairs_checked": int(len(s1)),
                         "level1_mismatches": int((r1 != g1[s1]).sum()), "bar": "bit-exact"}
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="C4", choices=list(SHAPES))
    ap.add_argument("--cpu-budget", type=float, default=12.0, help="seconds of CPU work for the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-level1", action="store_true", help="skip the level-1 P-matrix measurement")
    ap.add_argument("--level1-workload", default="C2", choices=["C1", "C2", "tiny"])
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--parity-rows", type=int, default=256, help="seeded random rows of every rank's slab checked against the oracle after the timed region")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    from sph_b200 import synth

    h, w, d, M, seed = SHAPES[args.workload]
    n = h * w
    k = K_NEIGHBOURS
    workload = f"{args.workload}: synthetic {'RGB image' if d == 3 else 'hyperspectral stack'} {h}x{w}x{d} (seed {seed}), exact kNN k={k} (L2, self included) + level-0 perplexity rows"
    config = {"workload": workload, "n_pixels": n, "channels": d, "k": k,
              "parallelism": f"query rows sharded over {world} GPU(s); database replicated by one NCCL all-gather of per-rank N/G-row uploads",
              "l2_flush": "inputs (operand matrices >= 2 GB at C4) are larger than L2" if n * d * 4 > 256e6 else "inputs fit L2; a 256 MB buffer is written between timed steps"}

    # ------------------------------------------------------------------ reference arm
    if args.impl == "reference":
        if rank != 0:
            return
        x = make_stack(synth, args.workload)
        orc = load_oracle()
        assert orc.blas_init()
        # a step = a bounded slab of query rows against the full database (the O(N^2 d) stage is linear in the query rows),
        # sized so that W + K steps end within a few minutes on 16 host cores
        slab = min(n, 2048 if n > 500000 else 4096)
        times, stats = [], []
        for it in range(args.warmup + args.steps):
            q0 = (it * slab) % max(1, n - slab)
            t0 = time.perf_counter()
            cpu_reference_step(orc, x, k, q0, q0 + slab)
            el = time.perf_counter() - t0
            if it >= args.warmup:
                times.append(el)
                stats.append(orc.knn_blas_last_stats())
        tot = sum(times)
        val = slab * len(times) / tot
        cores = os.cpu_count()
        sample = f"{slab} query rows against the full {n}-row database per step (faiss-style 4096x1024 blocked sgemm via OpenBLAS + per-row heaps, post-passes, Gaussian rows); CPU restatement (reference unbuildable here: faiss / HDILib / Eigen absent)"
        cb = {"value": val, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample}
        cb.update(cpu_sgemm_info(orc, d, stats))
        print(json.dumps({"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
                          "ms_per_step": 1e3 * tot / len(times), "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
                          "data": "synthetic", "config": config, "cpu_baseline": cb,
                          "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
        return

    # ------------------------------------------------------------------ our arm
    import torch

    from sph_b200 import capi, multigpu

    torch.cuda.set_device(local_rank)
    if world > 1:
        import torch.distributed as dist

        # stdout carries exactly one JSON line: NCCL prints its "NCCL version ..." banner on stdout when the communicator is
        # created (NCCL_DEBUG=VERSION/WARN), so file descriptor 1 points at stderr until the first collective is through
        with stdout_to_stderr():
            dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
            dist.barrier()
            torch.cuda.synchronize()
    capi.init(local_rank)

    x = make_stack(synth, args.workload)
    xh = torch.from_numpy(x).pin_memory()
    job = multigpu.ShardedKnnProw(n, d, k, rank, world)
    q0, q1 = job.q0, job.q1
    nq = q1 - q0
    job.replicate(xh)  # resident arm: the stack is in HBM before the timed region starts
    torch.cuda.synchronize()
    flush = torch.empty(64 << 20, dtype=torch.float32, device="cuda") if n * d * 4 <= 256e6 else None

    def step():
        if flush is not None:
            flush.fill_(1.0)
        job.compute(0.0)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def reduce_ranks(v: float, op="max") -> float:
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX if op == "max" else dist.ReduceOp.SUM)
        return float(t.item())

    for _ in range(args.warmup):
        step()
    barrier()
    capi.reset_launch_count()
    sweep_ms, prep_ms, rerank_ms, fallback_ms = [], [], [], []
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local_rank) as clocks:
        barrier()
        e0.record()
        for _ in range(args.steps):
            step()
            t = capi.knn_last_timing()
            sweep_ms.append(t["sweep_ms"]); prep_ms.append(t["prep_ms"]); rerank_ms.append(t["rerank_ms"]); fallback_ms.append(t["fallback_ms"])
        e1.record()
        barrier()
    launches = capi.launch_count()
    ms_total = reduce_ranks(e0.elapsed_time(e1))
    ms_per_step = ms_total / args.steps
    value = n * args.steps / (ms_total * 1e-3)
    stats = capi.knn_last_stats()

    # ---- parity of THIS configuration against the oracle (after the timed region; sampled rows of every rank's slab)
    parity = None
    orc = None
    if args.parity_rows > 0:
        orc = load_oracle(threads=max(1, (os.cpu_count() or 1) // world))
        rows, n_re, n_simt, tot_re, tot_simt = parity_rows(capi, n, q0, q1, args.parity_rows, 4242 + rank)
        sel = torch.from_numpy(rows - q0).cuda()
        got = (job.ids[sel].cpu().numpy(), job.dist[sel].cpu().numpy(), job.cols[sel].cpu().numpy().view(np.uint32), job.vals[sel].cpu().numpy())
        parity = parity_check(orc, x, k, rows, got)
        parity.update(resweep_rows_checked=n_re, simt_rows_checked=n_simt, resweep_rows_total=tot_re, simt_rows_total=tot_simt)

    # ---- roofline of the dominant kernel (the tcgen05 sweep): algorithmic flops = 2 * nq * n * d per launch
    peaks = _peaks()
    sweep = float(np.mean(sweep_ms))
    alg_flops = 2.0 * nq * n * d
    tensor_path = stats["tensor_core"] >= 1 and sweep > 0
    traffic = None
    tp = ROOT / "profiles" / "sweep_dram_bytes.json"
    if world == 1 and tp.exists():  # measured (ncu dram__bytes_read + write of the main sweep launch) for the single-GPU launch only
        try:
            traffic = json.loads(tp.read_text()).get(args.workload)
        except Exception:
            traffic = None
    if tensor_path:
        achieved = alg_flops / (sweep * 1e-3) / 1e12
        flavour = {4: "knn_sweep_pair", 3: "knn_sweep_resident<MH=2>", 2: "knn_sweep_resident<MH=1>", 1: "knn_sweep_stream"}.get(stats["tensor_core"])
        roofline = {"bound": "tensor", "kernel": f"{flavour}, {stats['passes']} fp16 product(s) per pair",
                    "achieved": achieved, "peak": peaks["tflops"], "unit": "TFLOP/s", "frac": achieved / peaks["tflops"], "traffic": traffic,
                    "peak_source": f"{peaks['src']} bf16 sustained (fp16 runs at the same tensor rate)",
                    "algorithmic_flops_per_launch": alg_flops,
                    "executed_multiplier": stats["kp"] / d, "executed_tflops": achieved * stats["kp"] / d,
                    "kernel_ms": sweep, "prep_ms": float(np.mean(prep_ms)), "rerank_ms": float(np.mean(rerank_ms)),
                    "fallback_ms": float(np.mean(fallback_ms)), "kernel_share_of_step": sweep / ms_per_step}
    else:
        # d < 8 (C3): the exact SIMT kernel, FP64-ALU bound: pairs/s against the FP32-lane issue peak
        pairs = float(nq) * n
        step_ms = e0.elapsed_time(e1) / args.steps
        sm_mhz = (clocks.summary().get("sm_mhz") or 1700.0)
        peak_pairs = 148 * 128 * sm_mhz * 1e6
        roofline = {"bound": "fp32-issue", "kernel": "knn_simt_kernel (fp32 screen + exact fp64 keys, warp per row)", "achieved": pairs / (step_ms * 1e-3),
                    "peak": peak_pairs, "unit": "pairs/s", "frac": pairs / (step_ms * 1e-3) / peak_pairs, "traffic": None,
                    "peak_source": "148 SMs x 128 lanes x sampled SM clock (one pair evaluation per lane-cycle)", "kernel_ms": step_ms}

    # ---- graph passes: Stage B2 (level-0 P rows) at the benchmarked size; the level-1 passes follow below
    s = torch.cuda.current_stream().cuda_stream
    p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    p0.record()
    for _ in range(5):
        capi.prow_gaussian_graph_dev(job.ids.data_ptr(), job.dist.data_ptr(), nq, k, 0.0, job.cols.data_ptr(), job.vals.data_ptr(), s)
    p1.record()
    torch.cuda.synchronize()
    prow_ms = p0.elapsed_time(p1) / 5
    prow_bytes = nq * (k * 12 + (k - 1) * 8)
    graph_passes = [{"pass": "B2 level-0 P rows", "kernel": "prow_graph_kernel (warp per row, fp32-screened fp64 bisection)", "bound": "hbm",
                     "achieved": prow_bytes / (prow_ms * 1e-3) / 1e9, "peak": peaks["hbm"], "unit": "GB/s",
                     "frac": prow_bytes / (prow_ms * 1e-3) / 1e9 / peaks["hbm"], "kernel_ms": prow_ms, "algorithmic_bytes_per_launch": prow_bytes,
                     "note": "instruction-issue bound, not HBM bound (ncu: issue slots 79 % busy, fp64 pipe 41 %, DRAM 2 %): ~20 bisection steps x k exponentials per row, DESIGN.md 4.2"}]

    # ---- end to end through the public API with HOST buffers: H2D of the stack + D2H of the four results inside the timed region
    out_h = (torch.empty((nq, k), dtype=torch.int64).pin_memory(), torch.empty((nq, k), dtype=torch.float32).pin_memory(),
             torch.empty((nq, k - 1), dtype=torch.int32).pin_memory(), torch.empty((nq, k - 1), dtype=torch.float32).pin_memory())
    e2e_steps = max(1, args.e2e_steps)
    if world == 1:
        # single GPU: the reference-facing C-ABI call sphb_knn_graph_prow (host pointers in, host pointers out)
        resident_ids = job.ids[:4096].cpu().numpy() if parity is not None else None
        del job
        torch.cuda.empty_cache()
        out_np = (out_h[0].numpy(), out_h[1].numpy(), out_h[2].numpy().view(np.uint32), out_h[3].numpy())
        es = torch.cuda.ExternalStream(capi.stream_ptr())
        capi.knn_graph_prow(xh.numpy(), k, 0.0, q0=q0, q1=q1, out=out_np)  # warm-up (workspace cache)
        barrier()
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record(es)
        for _ in range(e2e_steps):
            capi.knn_graph_prow(xh.numpy(), k, 0.0, q0=q0, q1=q1, out=out_np)
        f1.record(es)
        barrier()
        e2e_api = "sphb_knn_graph_prow (C ABI, host pointers)"
        e2e_same = bool(np.array_equal(out_np[0][:4096], resident_ids)) if resident_ids is not None else None
    else:
        # one process per GPU: sph_b200.multigpu.ShardedKnnProw — per-rank N/G-row upload, NCCL all-gather, slab search, per-rank download
        job(xh, out_h)  # warm-up
        barrier()
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record()
        for _ in range(e2e_steps):
            job(xh, out_h)
        f1.record()
        barrier()
        e2e_api = "sph_b200.multigpu.ShardedKnnProw (per-rank slab upload + ncclAllGather + C-ABI _dev calls + per-rank download)"
        e2e_same = bool(np.array_equal(out_h[0][:4096].numpy(), job.ids[:4096].cpu().numpy()))
    e2e_ms = reduce_ranks(f0.elapsed_time(f1))
    e2e_value = n * e2e_steps / (e2e_ms * 1e-3)
    h2d = n * d * 4
    d2h = n * (k * 12 + (k - 1) * 8)
    checksum = int(out_h[0][:, 1].sum().item() % (1 << 31))

    if parity is not None and world > 1:
        for key in ("rows_checked", "id_mismatch_rows", "dist_mismatch", "prow_col_mismatch_rows", "resweep_rows_checked", "simt_rows_checked",
                    "resweep_rows_total", "simt_rows_total"):
            parity[key] = int(reduce_ranks(float(parity[key]), "sum"))
        parity["prow_max_rel"] = reduce_ranks(parity["prow_max_rel"])
        parity["oracle_seconds"] = reduce_ranks(parity["oracle_seconds"])
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return
    if parity is not None:
        parity["checked_against"] = "oracle/oracle_knn.cpp orc_knn_exact_rows + orc_knn_postprocess_rows, oracle/oracle_prow.cpp orc_prow_gaussian_graph"
        parity["e2e_equals_resident_first_rows"] = e2e_same

    level1 = None
    if world == 1 and not args.no_level1:
        level1, passes = level1_p_matrix(capi, synth, orc if not args.no_cpu_baseline else None, peaks, args.level1_workload,
                                         cpu=not args.no_cpu_baseline and orc is not None)
        graph_passes += passes

    geodesic = None
    if world == 1 and not args.no_level1:
        try:
            geodesic = geodesic_block(capi, synth, orc if not args.no_cpu_baseline else None, args.level1_workload)
        except Exception as e:  # a side measurement must not take the bench line down
            geodesic = {"error": repr(e)[:300]}

    cpu_baseline = None
    if world == 1 and not args.no_cpu_baseline:
        orc = orc or load_oracle()
        slab = min(n, 2048 if n > 500000 else 4096)
        done, el, cstats = run_cpu_sample(orc, x, k, args.cpu_budget, slab)
        cpu_baseline = {"value": done / el, "unit": UNIT, "cores": os.cpu_count(), "kind": "port",
                        "sample": f"{done} query rows against the full {n}-row database in {el:.1f} s (faiss-style blocked sgemm via OpenBLAS + heaps, post-passes, Gaussian rows)"}
        cpu_baseline.update(cpu_sgemm_info(orc, d, cstats))

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f16 tensor-core filter (fp32 accumulate) + f64 exact re-rank",
            "data": "synthetic", "config": config, "clocks": clocks.summary(),
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "ms_per_step": e2e_ms / e2e_steps,
                    "steps": e2e_steps, "api": e2e_api, "host_buffers": "pinned caller buffers (the reference's are pageable std::vectors)",
                    "result_checksum": checksum},
            "gpu_launches": launches, "roofline": roofline, "graph_pass_roofline": graph_passes, "parity": parity, "level1_p_matrix": level1,
            "geodesic_distances": geodesic, "cpu_baseline": cpu_baseline, "knn_stats": stats}
    print(json.dumps(line))


if __name__ == "__main__":
    main()
