"""Writes the fixtures that pin code against THIRD-PARTY libraries the reference depends on but does not vendor:

  lz4_blocks_liblz4.json   LZ4 blocks produced by the real liblz4 (LZ4_compress_default, reached through pyarrow's codec "lz4_raw",
                           which bundles liblz4) for four payloads, plus one complete cache file in the layout of
                           utils::writeCompressedVecToBinary (sph/utils/FileIO.hpp:165-205) — what sph_b200/csrc/cache_io.cu must read
  geodesic_scipy_60.json   a 60-point kNN graph (true L2 edge weights) and scipy.sparse.csgraph.dijkstra's distances from three
                           sources — an independent Dijkstra for oracle/oracle_geodesic.cpp (Boost.Graph itself is absent)

Run here (needs pyarrow and scipy); the tests only read the JSON files:  python tests/golden/make_external_goldens.py"""
import base64
import hashlib
import json
import struct
import sys
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
sys.path.insert(0, str(HERE.parent))


def payloads():
    rng = np.random.default_rng(20261105)
    return {
        "int64_ramp_3000": np.arange(3000, dtype=np.int64).tobytes(),
        "float32_repeats": np.repeat(rng.random(64).astype(np.float32), 23).tobytes(),
        "random_bytes_5000": bytes(rng.integers(0, 256, 5000, dtype=np.uint8)),
        "zeros_40000": b"\x00" * 40000,
    }


def main():
    import pyarrow as pa
    from scipy.sparse import csr_matrix
    from scipy.sparse.csgraph import dijkstra

    from test_oracle_geodesic import knn_graph_l2

    codec = pa.Codec("lz4_raw")
    blocks = {}
    for name, data in payloads().items():
        blk = codec.compress(data, asbytes=True)
        blocks[name] = {"size": len(data), "sha256": hashlib.sha256(data).hexdigest(), "block_b64": base64.b64encode(blk).decode()}
    vec = (np.arange(2500, dtype=np.int64) * 7) % 1000
    raw = vec.tobytes()
    blk = codec.compress(raw, asbytes=True)
    file_bytes = struct.pack("<QQQ", len(raw), vec.size, 1) + struct.pack("<i", len(blk)) + blk
    out = {"source": f"liblz4 via pyarrow {pa.__version__} codec lz4_raw (LZ4 block format)", "blocks": blocks,
           "vec_file": {"dtype": "int64", "formula": "(arange(2500) * 7) % 1000", "file_b64": base64.b64encode(file_bytes).decode()}}
    (HERE / "lz4_blocks_liblz4.json").write_text(json.dumps(out, indent=1))

    rng = np.random.default_rng(60)
    x = rng.normal(size=(60, 4)).astype(np.float32)
    ids, dist = knn_graph_l2(x, 5)
    n, k = ids.shape
    rows = np.repeat(np.arange(n), k)
    keep = ids.reshape(-1) != rows
    g = csr_matrix((dist.reshape(-1)[keep].astype(np.float64), (rows[keep], ids.reshape(-1)[keep])), shape=(n, n))
    src = [0, 17, 59]
    d = dijkstra(g, directed=True, indices=src)
    geo = {"source": "scipy.sparse.csgraph.dijkstra (directed, fp64) on the fp32 edge weights below", "n": n, "k": k, "ids": ids.tolist(),
           "dist_f32_hex": [[float(v).hex() for v in r] for r in dist.astype(np.float64)], "sources": src,
           "distances": [[None if not np.isfinite(v) else float(v) for v in r] for r in d]}
    (HERE / "geodesic_scipy_60.json").write_text(json.dumps(geo))
    print("written")


if __name__ == "__main__":
    main()
