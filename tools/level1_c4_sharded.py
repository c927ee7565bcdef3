"""The north_star target on several GPUs (one process per GPU): exact kNN graph + level-0 P rows + LEVEL-1 similarities and P matrix
of a 2048 x 2048 x 128 stack (C4; level 1 = 2 x 2 pixel blocks = 1 048 576 components).

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 \
      tools/level1_c4_sharded.py --workload C4 --out gpurun_out/level1_c4_g8.json

Stages (time = wall clock between device synchronisations, max over ranks):
  A + B2   multigpu.ShardedKnnProw: query rows sharded, database replicated by one all-gather            
  gather   the level-0 P rows (cols, vals: n * (k - 1) * 8 bytes) all-gathered: the walks need the whole matrix
  RW1+RW2  bands of image rows (an even number: a 2 x 2 parent never straddles a band) dealt to the ranks round-robin; every band
           is walked (sphb_random_walks_dev on the band's start points) and merged (sphb_rw_merge_dev) where it is walked —
           a band's merged rows depend on that band's walks only (tools/level1_c4.py)
  gather   the merged rows of all bands, padded all-gather, reassembled in parent order on every rank
  D1..B6   multigpu.ShardedLevel: createSimilarities by output row, one gather of the slabs, level tail on every rank   (8e row D1)
Parity: --parity-rows rows of createSimilarities recomputed by the oracle from the same merged matrix on rank 0 (as level1_c4.py).

--emulate runs the same flow on gloo with CPU tensors against the CPU emulation build of the kernels (tests/emu; the kNN search
is then a numpy brute force, the tensor-core kernels cannot be emulated) — a logic check on a tiny stack:
  SPHB_EMULATE=1 SPHB_LIB=$PWD/tests/emu/_build/libsphb200_emu.so python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 \
      --master-addr 127.0.0.1 --master-port 29513 tools/level1_c4_sharded.py --emulate --shape 12,16,6,3,5 --k 7 --band-rows 4 --walks 6,3

NOT RUN ON HARDWARE YET (written after the GPU budget of round 2 was spent); the emulated 2- and 3-rank runs reproduce the
single-process result bit for bit."""
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
    ap.add_argument("--shape", default="", help="h,w,d,endmembers,seed instead of a named workload")
    ap.add_argument("--k", type=int, default=0)
    ap.add_argument("--band-rows", type=int, default=64)
    ap.add_argument("--walks", default="90,15", help="W,L")
    ap.add_argument("--parity-rows", type=int, default=128)
    ap.add_argument("--emulate", action="store_true")
    ap.add_argument("--out", default="")
    args = ap.parse_args()

    import torch
    import torch.distributed as dist

    import bench
    from sph_b200 import capi, host, synth
    from sph_b200.multigpu import ShardedKnnProw, ShardedLevel, slab_bounds

    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    emu = args.emulate
    if emu:
        assert capi.load().sphb_version() == -1, "--emulate needs the tests/emu build (SPHB_LIB)"
        dev = torch.device("cpu")
        if world > 1:
            dist.init_process_group("gloo")
        capi.init(0)
    else:
        torch.cuda.set_device(local)
        dev = torch.device("cuda", local)
        if world > 1:
            with bench.stdout_to_stderr():
                dist.init_process_group("nccl", device_id=dev)
                dist.barrier()
                torch.cuda.synchronize()
        capi.init(local)
    stream = 0 if emu else torch.cuda.current_stream().cuda_stream

    class Level(ShardedLevel):  # buffer placement: host memory with the emulation build
        def _dev(self):
            return dev

        def _stream(self, d):
            return stream

    def ptr(t):
        return t.data_ptr()

    def sync():
        if not emu:
            torch.cuda.synchronize()

    class Timer:
        def __init__(self):
            self.ms = {}

        def run(self, name, fn):
            if world > 1:
                dist.barrier()
            sync()
            t0 = time.perf_counter()
            r = fn()
            sync()
            t = torch.tensor([1e3 * (time.perf_counter() - t0)], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            self.ms[name] = self.ms.get(name, 0.0) + float(t)
            return r

    tm = Timer()
    if args.shape:
        bench.SHAPES["custom"] = tuple(int(v) for v in args.shape.split(","))
        args.workload = "custom"
    h, w, d, M, seed = bench.SHAPES[args.workload]
    k = args.k or bench.K_NEIGHBOURS
    W, L = (int(v) for v in args.walks.split(","))
    n = h * w
    assert args.band_rows % 2 == 0 and h % 2 == 0 and w % 2 == 0
    x = bench.make_stack(synth, args.workload)

    # ---- Stage A + B2, then the whole level-0 P matrix on every rank
    q0, q1, chunk = slab_bounds(n, rank, world)
    cols_all = torch.zeros((world * chunk, k - 1), dtype=torch.int32, device=dev)
    vals_all = torch.zeros((world * chunk, k - 1), dtype=torch.float32, device=dev)
    if emu:  # the tensor-core search cannot be emulated: numpy brute force for the graph, the emulated kernel for the P rows
        from test_oracle_geodesic import knn_graph_l2

        ids_np, dist_np = knn_graph_l2(x, k)
        ids_t = torch.from_numpy(ids_np[q0:q1].copy())
        dist_t = torch.from_numpy(dist_np[q0:q1].copy())

        def stage_a():
            if q1 > q0:
                capi.prow_gaussian_graph_dev(ptr(ids_t), ptr(dist_t), q1 - q0, k, 0.0, ptr(cols_all[q0:]), ptr(vals_all[q0:]), stream)
    else:
        job = ShardedKnnProw(n, d, k, rank, world)
        xh = torch.from_numpy(x).pin_memory()

        def stage_a():
            job.replicate(xh)
            job.compute(0.0)
            if q1 > q0:
                cols_all[q0:q1].copy_(job.cols)
                vals_all[q0:q1].copy_(job.vals)
    tm.run("knn_prow", stage_a)

    def gather_p():
        if world > 1:
            for buf in (cols_all, vals_all):
                dist.all_gather_into_tensor(buf, buf[rank * chunk:(rank + 1) * chunk].clone())
    tm.run("gather_level0_p", gather_p)
    if not emu:
        del job, xh
        torch.cuda.empty_cache()
    del x

    # ---- RW1 + RW2 per band; the bands are dealt to the ranks round-robin
    hb, wb = h // 2, w // 2
    n1 = hb * wb
    rows, colsi = np.divmod(np.arange(n), w)
    parents = ((rows // 2) * wb + colsi // 2).astype(np.int64)
    dpar = torch.from_numpy(parents).to(dev)
    rp0 = torch.arange(n + 1, dtype=torch.int64, device=dev) * (k - 1)
    dc0, dv0 = cols_all[:n].reshape(-1), vals_all[:n].reshape(-1)
    _, ks = host.level_schedule([n, n1], k)
    k1 = int(ks[1])
    band = args.band_rows * w
    bands = [(b0, min(n, b0 + band)) for b0 in range(0, n, band)]
    cap_w = band * W * L
    wrp_b = torch.empty(band + 1, dtype=torch.int64, device=dev)
    wc = torch.empty(cap_w, dtype=torch.int32, device=dev)
    wv = torch.empty(cap_w, dtype=torch.float32, device=dev)
    wrp_full = torch.empty(n + 1, dtype=torch.int64, device=dev)
    mrp_b = torch.empty(n1 + 1, dtype=torch.int64, device=dev)
    mc_b = torch.empty(cap_w, dtype=torch.int32, device=dev)
    mv_b = torch.empty(cap_w, dtype=torch.float32, device=dev)
    mine = []  # (first parent, counts per parent int64, cols, vals) of my bands

    def walk_bands():
        for bi, (b0, b1) in enumerate(bands):
            if bi % world != rank:
                continue
            wn = capi.random_walks_dev(n, ptr(rp0), ptr(dc0), ptr(dv0), W, L, capi.W_NORMAL, 1, b0, b1, ptr(wrp_b), ptr(wc), ptr(wv), cap_w, stream)
            wrp_full[: b0 + 1] = 0
            wrp_full[b0: b1 + 1] = wrp_b[: b1 - b0 + 1]
            wrp_full[b1 + 1:] = wn
            mn = capi.rw_merge_dev(n, ptr(wrp_full), ptr(wc), ptr(wv), wn, ptr(dpar), n1, True, True, ptr(mrp_b), ptr(mc_b), ptr(mv_b), cap_w, stream)
            p0, p1 = (b0 // w // 2) * wb, ((b1 // w + 1) // 2) * wb
            sync()
            assert int(mrp_b[p0]) == 0 and int(mrp_b[p1]) == mn, "a band's merged rows must lie inside the band's parents"
            mine.append((p0, (mrp_b[p0 + 1: p1 + 1] - mrp_b[p0: p1]).clone(), mc_b[:mn].clone(), mv_b[:mn].clone()))
    tm.run("walks_and_merge", walk_bands)
    del wc, wv, mc_b, mv_b, wrp_full, wrp_b, mrp_b, rp0

    # ---- the merged matrix on every rank: counts first, then the padded entry slabs
    def gather_merged():
        counts = torch.zeros(n1, dtype=torch.int64, device=dev)
        for p0, cnt, _, _ in mine:
            counts[p0: p0 + cnt.numel()] = cnt
        if world > 1:
            dist.all_reduce(counts)  # every parent belongs to exactly one band, so the sum is a scatter
        mrp = torch.zeros(n1 + 1, dtype=torch.int64, device=dev)
        mrp[1:] = torch.cumsum(counts, 0)
        total = int(mrp[-1])
        my_n = sum(int(c.numel()) for _, _, c, _ in mine)
        sizes = torch.tensor([my_n], dtype=torch.int64, device=dev)
        all_sizes = [torch.zeros_like(sizes) for _ in range(world)]
        if world > 1:
            dist.all_gather(all_sizes, sizes)
        else:
            all_sizes = [sizes]
        mx = max(max(int(s) for s in all_sizes), 1)
        mc = torch.empty(total, dtype=torch.int32, device=dev)
        mv = torch.empty(total, dtype=torch.float32, device=dev)
        for dtype, pick, outbuf in ((torch.int32, 2, mc), (torch.float32, 3, mv)):
            sendbuf = torch.zeros(mx, dtype=dtype, device=dev)
            at = 0
            for item in mine:
                sendbuf[at: at + item[pick].numel()] = item[pick]
                at += item[pick].numel()
            parts = [torch.empty_like(sendbuf) for _ in range(world)]
            if world > 1:
                dist.all_gather(parts, sendbuf)
            else:
                parts = [sendbuf]
            for r in range(world):  # rank r's slab holds its bands in band order
                at = 0
                for bi, (b0, b1) in enumerate(bands):
                    if bi % world != r:
                        continue
                    p0, p1 = (b0 // w // 2) * wb, ((b1 // w + 1) // 2) * wb
                    e0, e1 = int(mrp[p0]), int(mrp[p1])
                    outbuf[e0:e1] = parts[r][at: at + (e1 - e0)]
                    at += e1 - e0
        return mrp, mc, mv
    mrp, mc, mv = tm.run("gather_merged", gather_merged)
    mine.clear()
    merged = int(mc.numel())

    # ---- the level: D1 by output row over the ranks, gather, tail on every rank
    lvl = Level(rank, world)
    prp, pc, pv = tm.run("level1_prob_dist", lambda: lvl(mrp, mc, mv, k1, 0.0))
    res = {"workload": f"{args.workload}: {h}x{w}x{d}, k = {k}; level 1 = 2x2 pixel blocks ({n1} components), k_1 = {k1}, walks W={W} L={L} NORMAL, "
                       f"bands of {args.band_rows} image rows dealt round-robin",
           "n_gpus": world, "emulated": emu, "stages_ms": tm.ms, "total_ms": sum(tm.ms.values()), "merged_nnz": merged, "p_nnz": int(pc.numel())}
    # every rank holds the same P matrix
    chk = torch.tensor([float(pv.double().sum()), float(pc.long().sum()), float(prp[-1])], dtype=torch.float64, device=dev)
    lo, hi = chk.clone(), chk.clone()
    if world > 1:
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    res["all_ranks_hold_the_same_p_matrix"] = bool(torch.equal(lo, hi))
    res["p_checksum"] = [float(v) for v in chk]
    if rank == 0:
        if args.parity_rows > 0:
            import oracle_lib

            orc = oracle_lib.load()
            sc = torch.empty((args.parity_rows, k1), dtype=torch.int32, device=dev)
            sv = torch.empty((args.parity_rows, k1), dtype=torch.float32, device=dev)
            snz = torch.empty(args.parity_rows, dtype=torch.int32, device=dev)
            r0 = min(max(0, (n1 // 2 // wb) * wb + wb // 3), max(0, n1 - args.parity_rows))
            r1 = min(n1, r0 + args.parity_rows)
            capi.create_similarities_slab_dev(n1, ptr(mrp), ptr(mc), ptr(mv), merged, k1, 1e-4, 0, None, r0, r1, ptr(sc), ptr(sv), ptr(snz), stream)
            sync()
            oc, ov, onz = orc.create_similarities_rows(mrp.cpu().numpy(), mc.cpu().numpy().view(np.uint32), mv.cpu().numpy(), k1, r0, r1, 1e-4, 0)
            gc, gv, gnz = sc.cpu().numpy().view(np.uint32), sv.cpu().numpy(), snz.cpu().numpy()
            same, worst = 0, 0.0
            for i in range(r1 - r0):
                m = int(onz[i])
                if gnz[i] == m and np.array_equal(gc[i, :m], oc[i, :m]):
                    same += 1
                    if m:
                        worst = max(worst, float((np.abs(gv[i, :m] - ov[i, :m]) / np.maximum(np.abs(ov[i, :m]), 1e-30)).max()))
            res["parity"] = {"rows_checked": int(r1 - r0), "rows_identical_pattern": int(same), "max_rel_on_identical_rows": worst}
        line = json.dumps(res)
        print(line)
        if args.out:
            with open(args.out, "w") as f:
                f.write(line + "\n")
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
