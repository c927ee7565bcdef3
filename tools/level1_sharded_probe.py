"""SURVEY 8(e), D1 sharded by output row — measurement probe for multigpu.ShardedLevel (one process per GPU):

  python -m torch.distributed.run --nnodes=1 --nproc-per-node G --master-addr 127.0.0.1 --master-port 29511 \
      tools/level1_sharded_probe.py --workload C2 --out gpurun_out/level1_sharded_gG.json

Rank 0 builds level 0 (kNN + P rows), the walks and the merged walk matrix of a 2 x 2 block hierarchy (as bench.py's level-1
block does); the matrix is broadcast (replicate_csr), then timed per rank with CUDA events, max over ranks:
  * the one-GPU level (sphb_level_prob_dist_dev) on rank 0,
  * ShardedLevel on all ranks (slab D1 + three all-gathers + the level tail on every rank),
and the sharded P matrix of every rank is compared with rank 0's one-GPU result (row pointers and columns exact, values to the
kernel flavour's summation order).

NOT RUN ON HARDWARE YET: written after the GPU budget of round 2 was spent; the pieces it calls are covered by
tests/test_zy_gpu_sims_slab.py (emulation) and tests/test_host.py (gloo)."""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="C2")
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--out", default="")
    args = ap.parse_args()

    import torch
    import torch.distributed as dist

    import bench
    from sph_b200 import capi, host, synth
    from sph_b200.multigpu import ShardedLevel

    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        with bench.stdout_to_stderr():  # NCCL prints its banner on stdout when the communicator is created
            dist.init_process_group("nccl", device_id=torch.device("cuda", local))
            dist.barrier()
            torch.cuda.synchronize()
    capi.init(local)
    s = torch.cuda.current_stream().cuda_stream
    h, w, d, M, seed = bench.SHAPES[args.workload]
    k = bench.K_NEIGHBOURS
    n = h * w
    hb, wb = (h + 1) // 2, (w + 1) // 2
    n1 = hb * wb
    _, ks = host.level_schedule([n, n1], k)
    k1 = int(ks[1])
    mrp = mc = mv = None
    if rank == 0:  # level 0, walks, merge: once
        x = bench.make_stack(synth, args.workload)
        ids, dst, cols, vals = capi.knn_graph_prow(x, k, 0.0)
        del x, ids, dst
        W, L = 90, 15
        rp0 = torch.arange(n + 1, dtype=torch.int64, device="cuda") * (k - 1)
        dc0 = torch.from_numpy(cols.reshape(-1).view(np.int32)).cuda()
        dv0 = torch.from_numpy(vals.reshape(-1)).cuda()
        cap_w = n * W * L // 2
        wrp = torch.empty(n + 1, dtype=torch.int64, device="cuda")
        wc = torch.empty(cap_w, dtype=torch.int32, device="cuda")
        wv = torch.empty(cap_w, dtype=torch.float32, device="cuda")
        wnnz = capi.random_walks_dev(n, rp0.data_ptr(), dc0.data_ptr(), dv0.data_ptr(), W, L, capi.W_NORMAL, 1, 0, n, wrp.data_ptr(), wc.data_ptr(),
                                     wv.data_ptr(), cap_w, s)
        rows, colsi = np.divmod(np.arange(n), w)
        parents = ((rows // 2) * wb + colsi // 2).astype(np.uint64)
        dpar = torch.from_numpy(parents.view(np.int64)).cuda()
        mrp = torch.empty(n1 + 1, dtype=torch.int64, device="cuda")
        mc = torch.empty(wnnz, dtype=torch.int32, device="cuda")
        mv = torch.empty(wnnz, dtype=torch.float32, device="cuda")
        mnnz = capi.rw_merge_dev(n, wrp.data_ptr(), wc.data_ptr(), wv.data_ptr(), wnnz, dpar.data_ptr(), n1, True, True, mrp.data_ptr(), mc.data_ptr(),
                                 mv.data_ptr(), wnnz, s)
        mc, mv = mc[:mnnz].contiguous(), mv[:mnnz].contiguous()
        del wrp, wc, wv, rp0, dc0, dv0
    job = ShardedLevel(rank, world)
    mrp, mc, mv = job.replicate_csr(mrp, mc, mv, src=0)

    def timed(fn):
        out = fn()  # warm-up
        best = None
        for _ in range(args.reps):
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            out = fn()
            e1.record()
            torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1)], device="cuda")
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            best = float(t) if best is None else min(best, float(t))
        return best, out

    sh_ms, (srp, sc, sv) = timed(lambda: job(mrp, mc, mv, k1, 0.0))
    res = {"workload": args.workload, "n_gpus": world, "components": n1, "k1": k1, "merged_nnz": int(mc.numel()), "sharded_level_ms": sh_ms}
    # the one-GPU level on rank 0, and every rank's sharded result against it
    cap_p = 2 * n1 * k1
    prp = torch.empty(n1 + 1, dtype=torch.int64, device="cuda")
    pc = torch.empty(cap_p, dtype=torch.int32, device="cuda")
    pv = torch.empty(cap_p, dtype=torch.float32, device="cuda")
    one_ms, pnnz = timed(lambda: capi.level_prob_dist_dev(n1, mrp.data_ptr(), mc.data_ptr(), mv.data_ptr(), mc.numel(), k1, 1e-4, 0, None, True, 0.0, 1, True,
                                                          prp.data_ptr(), pc.data_ptr(), pv.data_ptr(), cap_p, s))
    same = bool(torch.equal(srp, prp) and sc.numel() == pnnz and torch.equal(sc, pc[:pnnz]) and torch.allclose(sv, pv[:pnnz], rtol=2e-5, atol=1e-7))
    flag = torch.tensor([1 if same else 0], device="cuda")
    if world > 1:
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    res.update({"one_gpu_level_ms": one_ms, "speedup": one_ms / sh_ms, "all_ranks_equal_one_gpu_result": bool(int(flag))})
    if rank == 0:
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
