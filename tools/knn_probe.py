"""Ad-hoc timing probe of the device-resident kNN path (not the bench)."""
import sys, time
import numpy as np, torch
sys.path.insert(0, '.')
from sph_b200 import capi, synth

h, w, d, k = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
nq = int(sys.argv[5]) if len(sys.argv) > 5 else h * w
capi.init()
t0 = time.time()
x = synth.hyperspectral(h, w, d, endmembers=32, seed=1004)
print("gen", time.time() - t0, flush=True)
n = h * w
dx = torch.from_numpy(x).cuda()
ids = torch.empty((nq, k), dtype=torch.int64, device="cuda")
dist = torch.empty((nq, k), dtype=torch.float32, device="cuda")
s = torch.cuda.current_stream().cuda_stream
import os
for it in range(int(os.environ.get("PROBE_ITERS", "3"))):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    capi.knn_exact_dev(dx.data_ptr(), n, d, k, capi.METRIC_L2, False, 0, nq, ids.data_ptr(), dist.data_ptr(), s)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    st = capi.knn_last_stats()
    fl = 2.0 * nq * n * d
    print(f"iter {it}: {ms:.1f} ms  alg {fl/ms/1e9:.1f} TF/s  exec {2.0*nq*n*st['kp']/ms/1e9:.1f} TF/s  stats {st}  accepted/row {st['accepted']/nq:.0f} compactions/row {st['compactions']/nq:.2f} timing {capi.knn_last_timing()}", flush=True)
