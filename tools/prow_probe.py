"""P-row kernel probe: n rows of k sorted synthetic distances through sphb_prow_gaussian_graph_dev (device resident).
argv: n k [scale].  Prints ms per call and GB/s of the 1812-B-per-row model."""
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from sph_b200 import capi  # noqa: E402

n, k = int(sys.argv[1]), int(sys.argv[2])
scale = float(sys.argv[3]) if len(sys.argv) > 3 else 0.3
capi.init(0)
g = torch.Generator(device="cuda").manual_seed(5)
dist = torch.sort(torch.rand((n, k), generator=g, device="cuda") * scale + 0.05, dim=1).values.float()
dist[:, 0] = 0
ids = torch.randint(0, n, (n, k), generator=g, device="cuda", dtype=torch.int64)
cols = torch.empty((n, k - 1), dtype=torch.int32, device="cuda")
vals = torch.empty((n, k - 1), dtype=torch.float32, device="cuda")
s = torch.cuda.current_stream().cuda_stream
for it in range(3):
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    capi.prow_gaussian_graph_dev(ids.data_ptr(), dist.data_ptr(), n, k, 0.0, cols.data_ptr(), vals.data_ptr(), s)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    print(f"prow n={n} k={k}: {ms:.3f} ms, {n * (k * 12 + (k - 1) * 8) / ms / 1e6:.1f} GB/s", flush=True)
