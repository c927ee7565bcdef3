"""Single-process multi-device entry point at bench size: sphb_knn_graph_prow_multi on all visible GPUs from one host thread
(C4 by default), wall clock around the call (host buffers in, host buffers out: an e2e number)."""
import json
import sys
import time
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import bench  # noqa: E402
from sph_b200 import capi, synth  # noqa: E402


def main():
    import torch

    workload = sys.argv[1] if len(sys.argv) > 1 else "C4"
    g = torch.cuda.device_count()
    x = bench.make_stack(synth, workload)
    n = x.shape[0]
    k = bench.K_NEIGHBOURS
    out = tuple(torch.empty(s, dtype=t).pin_memory().numpy() for s, t in (((n, k), torch.int64), ((n, k), torch.float32), ((n, k - 1), torch.int32),
                                                                          ((n, k - 1), torch.float32)))
    out = (out[0], out[1], out[2].view(np.uint32), out[3])
    xp = torch.from_numpy(x).pin_memory().numpy()
    times = []
    for _ in range(3):
        t0 = time.perf_counter()
        capi.knn_graph_prow_multi(xp, k, list(range(g)), out=out)
        times.append(time.perf_counter() - t0)
    print(json.dumps({"workload": workload, "devices": g, "seconds": times, "pixels_per_s_best": n / min(times),
                      "api": "sphb_knn_graph_prow_multi (one host thread, worker thread per device, peer copies)",
                      "checksum": int(out[0][::4099, 1].sum())}))


if __name__ == "__main__":
    main()
