#!/usr/bin/env python
"""bench.py — headline benchmark of the SPH hot path on B200.

A "step" is one pass of the hot path over one synthetic image stack:
exact kNN graph (Stage A, tcgen05 sweep + exact re-rank) followed by the level-0
perplexity rows (Stage B) for every pixel; query rows are sharded over the ranks
(database replicated, no data-path collective), so `scaling` is "strong".

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload C4] [--impl reference]

Prints ONE JSON line (rank 0).  See DESIGN.md §Measurement for every field.
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


def cpu_reference_step(orc, x, k, q0, q1):
    """The reference's CPU path for a slab of query rows: faiss-style blocked sgemm + heaps
    (oracle/oracle_knn.cpp orc_knn_blas_l2), post-passes, level-0 Gaussian rows."""
    ids, dist = orc.knn_blas_l2(x, k, q0, q1)
    orc.knn_postprocess(ids, dist, False, np.arange(q0, q1))
    orc.prow_gaussian_graph(ids, dist, 0.0)
    return ids, dist


def run_cpu_sample(x, k, budget_s, slab):
    sys.path.insert(0, str(ROOT / "tests"))
    import oracle_lib

    orc = oracle_lib.load()
    if not orc.blas_init():
        raise RuntimeError("OpenBLAS (numpy.libs) not found for the CPU baseline")
    n = x.shape[0]
    done = 0
    t0 = time.perf_counter()
    q = 0
    while True:
        q1 = min(n, q + slab)
        cpu_reference_step(orc, x, k, q, q1)
        done += q1 - q
        q = q1 % n
        el = time.perf_counter() - t0
        if el >= budget_s or done >= n:
            break
    return done, el


def level1_p_wall_time(capi, synth):
    """Second half of BASELINE.json's metric: wall time of the level-1 P matrix on a FIXED hierarchy.
    Salinas-shaped stack (C2): level-0 kNN + P rows and W=90 x L=15 random walks are inputs (not timed); level 1 = 2x2 pixel
    blocks (the ImageHierarchy merge loop is host code and out of scope).  Timed, through the host-buffer C ABI:
    mergeNodesRandomWalks -> createSimilarities(k_1, 1e-4) -> in-place Gaussian rows (ignore = -1) -> symmetrizeTSNE
    (sph/LevelSimilarities.cpp:460-508, 589-623)."""
    from sph_b200 import host

    h, w, d, k = 512, 217, 204, 91
    x = synth.hyperspectral(h, w, d, endmembers=16, seed=1002)
    n = h * w
    ids, dist, cols, vals = capi.knn_graph_prow(x, k, 0.0)
    rp0 = np.arange(n + 1, dtype=np.int64) * (k - 1)
    t0 = time.perf_counter()
    wrp, wc, wv = capi.random_walks(rp0, cols.reshape(-1), vals.reshape(-1), 90, 15, capi.W_NORMAL, seed=1, cap=n * 90 * 15 // 2)
    t_walks = time.perf_counter() - t0
    rows, colsi = np.divmod(np.arange(n), w)
    hb, wb = (h + 1) // 2, (w + 1) // 2
    parents = ((rows // 2) * wb + colsi // 2).astype(np.uint64)
    n1 = hb * wb
    _, ks = host.level_schedule([n, n1], k)
    k1 = ks[1]
    times = []
    for _ in range(3):
        t0 = time.perf_counter()
        mrp, mc, mv = capi.rw_merge(wrp, wc, wv, parents, n1, norm=True, weight_by_size=True)
        sc, sv, snz = capi.create_similarities(mrp, mc, mv, k1, 1e-4)
        prp = np.concatenate([[0], np.cumsum(snz, dtype=np.int64)])
        keep = np.arange(sc.shape[1])[None, :] < snz[:, None]
        pc, pv = sc[keep], sv[keep]
        pv = capi.prow_gaussian_inplace(prp, pv, 0.0, -1)
        orp, oc, ov = capi.symmetrize(prp, pc, pv)
        times.append(time.perf_counter() - t0)
    return {"workload": f"C2-shaped {h}x{w}x{d}: level 1 = 2x2 pixel blocks ({n1} components), k_1 = {k1}, walks W=90 L=15 NORMAL",
            "wall_ms": 1e3 * min(times), "wall_ms_all": [1e3 * t for t in times], "random_walks_level0_wall_ms": 1e3 * t_walks,
            "walk_nnz": int(wrp[-1]), "merged_nnz": int(mrp[-1]), "p_nnz": int(orp[-1])}


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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="C4", choices=["C1", "C2", "C4", "C5", "tiny"])
    ap.add_argument("--cpu-budget", type=float, default=12.0, help="seconds of CPU work for the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-level1", action="store_true", help="skip the level-1 P-matrix wall-time measurement")
    ap.add_argument("--e2e-steps", type=int, default=1)
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    from sph_b200 import synth

    shapes = {"C1": (145, 145, 200, 16, 1001), "C2": (512, 217, 204, 16, 1002), "C4": (2048, 2048, 128, 32, 1004), "C5": (1024, 1024, 1000, 64, 1005),
              "tiny": (128, 128, 128, 8, 7)}
    h, w, d, M, seed = shapes[args.workload]
    n = h * w
    k = K_NEIGHBOURS
    workload = f"{args.workload}: synthetic hyperspectral stack {h}x{w}x{d} (seed {seed}), exact kNN k={k} (L2, self included) + level-0 perplexity rows"
    config = {"workload": workload, "n_pixels": n, "channels": d, "k": k, "parallelism": f"query rows sharded over {world} GPU(s), database replicated",
              "l2_flush": "inputs (operand matrices >= 2 GB) are larger than L2"}

    # ------------------------------------------------------------------ reference arm
    if args.impl == "reference":
        if rank != 0:
            return
        x = synth.hyperspectral(h, w, d, endmembers=M, seed=seed)
        sys.path.insert(0, str(ROOT / "tests"))
        import oracle_lib

        orc = oracle_lib.load()
        assert orc.blas_init()
        slab = 2048 if n > 500000 else min(n, 4096)
        times = []
        for it in range(args.warmup + args.steps):
            q0 = (it * slab) % max(1, n - slab)
            t0 = time.perf_counter()
            cpu_reference_step(orc, x, k, q0, q0 + slab)
            el = time.perf_counter() - t0
            if it >= args.warmup:
                times.append(el)
        tot = sum(times)
        val = slab * len(times) / tot
        cores = os.cpu_count()
        sample = f"{slab} query rows against the full {n}-row database per step (faiss-style blocked sgemm + heaps, post-passes, Gaussian rows); CPU restatement (reference unbuildable here)"
        print(json.dumps({"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
                          "ms_per_step": 1e3 * tot / len(times), "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
                          "data": "synthetic", "config": config,
                          "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
                          "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
        return

    # ------------------------------------------------------------------ our arm
    import torch

    from sph_b200 import capi

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

    x = synth.hyperspectral(h, w, d, endmembers=M, seed=seed)
    q0 = (n * rank) // world
    q1 = (n * (rank + 1)) // world
    nq = q1 - q0

    dx = torch.from_numpy(x).cuda()
    ids = torch.empty((nq, k), dtype=torch.int64, device="cuda")
    dst = torch.empty((nq, k), dtype=torch.float32, device="cuda")
    cols = torch.empty((nq, k - 1), dtype=torch.int32, device="cuda")
    vals = torch.empty((nq, k - 1), dtype=torch.float32, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream

    def step():
        capi.knn_exact_dev(dx.data_ptr(), n, d, k, capi.METRIC_L2, False, q0, q1, ids.data_ptr(), dst.data_ptr(), stream)
        capi.prow_gaussian_graph_dev(ids.data_ptr(), dst.data_ptr(), nq, k, 0.0, cols.data_ptr(), vals.data_ptr(), stream)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v: float) -> float:
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
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
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    ms_per_step = ms_total / args.steps
    value = n * args.steps / (ms_total * 1e-3)
    stats = capi.knn_last_stats()

    # roofline of the dominant kernel (the tcgen05 sweep): algorithmic flops = 2 * nq * n * d per launch
    peaks = _peaks()
    sweep = float(np.mean(sweep_ms))
    alg_flops = 2.0 * nq * n * d
    achieved = alg_flops / (sweep * 1e-3) / 1e12
    traffic = None
    tp = ROOT / "profiles" / "sweep_dram_bytes.json"
    if tp.exists():
        try:
            traffic = json.loads(tp.read_text()).get(args.workload)
        except Exception:
            traffic = None
    flavour = {4: "knn_sweep_pair (cta_group::2, 2 x 256 resident query rows)", 3: "knn_sweep_resident<MH=2>", 2: "knn_sweep_resident<MH=1>", 1: "knn_sweep_stream"}.get(stats["tensor_core"], "knn_simt")
    roofline = {"bound": "tensor", "kernel": f"{flavour}, {stats['passes']} fp16 product(s) per pair",
                "achieved": achieved, "peak": peaks["tflops"], "unit": "TFLOP/s", "frac": achieved / peaks["tflops"], "traffic": traffic,
                "peak_source": f"{peaks['src']} bf16 sustained (fp16 runs at the same tensor rate)",
                "algorithmic_flops_per_launch": alg_flops,
                "executed_multiplier": stats["kp"] / d, "executed_tflops": achieved * stats["kp"] / d,
                "kernel_ms": sweep, "prep_ms": float(np.mean(prep_ms)), "rerank_ms": float(np.mean(rerank_ms)),
                "fallback_ms": float(np.mean(fallback_ms)), "kernel_share_of_step": sweep / ms_per_step}

    # graph pass (Stage B2, level-0 P rows): HBM-bound, algorithmic bytes per row = k*4 + k*8 read + (k-1)*8 written
    p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    p0.record()
    for _ in range(5):
        capi.prow_gaussian_graph_dev(ids.data_ptr(), dst.data_ptr(), nq, k, 0.0, cols.data_ptr(), vals.data_ptr(), stream)
    p1.record()
    torch.cuda.synchronize()
    prow_ms = p0.elapsed_time(p1) / 5
    prow_bytes = nq * (k * 12 + (k - 1) * 8)
    graph_pass = {"kernel": "prow_graph_kernel (perplexity rows, warp per row)", "bound": "hbm", "achieved": prow_bytes / (prow_ms * 1e-3) / 1e9,
                  "peak": peaks["hbm"], "unit": "GB/s", "frac": prow_bytes / (prow_ms * 1e-3) / 1e9 / peaks["hbm"], "kernel_ms": prow_ms,
                  "algorithmic_bytes_per_launch": prow_bytes,
                  "note": "reported against the HBM peak; the kernel is FP64-exp bound (about 20 bisection steps x k fp64 exp per row), DESIGN.md 4.2"}

    # end to end through the host-buffer C ABI (pinned host memory; H2D + D2H inside the timed region)
    xh = torch.from_numpy(x).pin_memory()
    out_h = (torch.empty((nq, k), dtype=torch.int64).pin_memory(), torch.empty((nq, k), dtype=torch.float32).pin_memory(),
             torch.empty((nq, k - 1), dtype=torch.int32).pin_memory(), torch.empty((nq, k - 1), dtype=torch.float32).pin_memory())
    out_np = (out_h[0].numpy(), out_h[1].numpy(), out_h[2].numpy().view(np.uint32), out_h[3].numpy())
    del dx, ids, dst, cols, vals
    torch.cuda.empty_cache()
    es = torch.cuda.ExternalStream(capi.stream_ptr())
    capi.knn_graph_prow(xh.numpy(), k, 0.0, q0=q0, q1=q1, out=out_np)  # warm-up (allocator pools)
    barrier()
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    f0.record(es)
    for _ in range(args.e2e_steps):
        capi.knn_graph_prow(xh.numpy(), k, 0.0, q0=q0, q1=q1, out=out_np)
    f1.record(es)
    barrier()
    e2e_ms = max_over_ranks(f0.elapsed_time(f1))
    e2e_steps = max(1, args.e2e_steps)
    e2e_value = n * args.e2e_steps / (e2e_ms * 1e-3) if args.e2e_steps else None
    h2d = world * n * d * 4
    d2h = n * (k * 12 + (k - 1) * 8)
    checksum = int(out_np[0][:, 1].sum() % (1 << 31))

    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return

    level1 = None
    if world == 1 and not args.no_level1:
        level1 = level1_p_wall_time(capi, synth)

    cpu_baseline = None
    if world == 1 and not args.no_cpu_baseline:
        slab = 1024 if n > 500000 else min(n, 2048)
        done, el = run_cpu_sample(x, k, args.cpu_budget, slab)
        cpu_baseline = {"value": done / el, "unit": UNIT, "cores": os.cpu_count(), "kind": "port",
                        "sample": f"{done} query rows against the full {n}-row database in {el:.1f} s (faiss-style blocked sgemm via OpenBLAS + heaps, post-passes, Gaussian rows)"}

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f16 tensor-core filter (fp32 accumulate) + f64 exact re-rank",
            "data": "synthetic", "config": config, "clocks": clocks.summary(),
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "ms_per_step": e2e_ms / e2e_steps,
                    "result_checksum": checksum},
            "gpu_launches": launches, "roofline": roofline, "graph_pass_roofline": graph_pass, "level1_p_matrix": level1, "cpu_baseline": cpu_baseline,
            "knn_stats": stats}
    print(json.dumps(line))


if __name__ == "__main__":
    main()
