"""SURVEY 8(f)-4, cache blobs (sph_b200/csrc/cache_io.cu): the reference's LZ4 stage-cache files.  Host code — runs without a GPU.
liblz4 itself is absent from the image as a C library, but pyarrow bundles it: codec "lz4_raw" is LZ4 block format
(LZ4_compress_default / LZ4_decompress_safe), so both directions are pinned against the real library:
  * blocks written here decode with liblz4,
  * files assembled here with liblz4-compressed blocks in the reference's layout (sph/utils/FileIO.hpp:165-291) load here."""
import struct

import numpy as np
import pyarrow as pa
import pytest

LZ4 = pa.Codec("lz4_raw")


def payloads():
    rng = np.random.default_rng(0)
    yield b""
    yield b"a"
    yield b"abcdefghijkl"             # below the 13-byte minimum for a match
    yield b"\x00" * 100_000            # one long run (overlapping match, extended lengths)
    yield bytes(rng.integers(0, 256, 70_000, dtype=np.uint8))   # incompressible
    yield np.arange(50_000, dtype=np.int64).tobytes()            # kNN-index-like
    yield np.repeat(rng.random(300).astype(np.float32), 17).tobytes()
    yield (b"0123456789abcdef" * 5000) + bytes(rng.integers(0, 256, 333, dtype=np.uint8))


def test_blocks_decode_with_liblz4_and_back(capi):
    for data in payloads():
        blk = capi.lz4_compress(data)
        assert len(blk) <= len(data) + len(data) // 255 + 16
        assert LZ4.decompress(blk, decompressed_size=len(data), asbytes=True) == data if data else blk == b"\x00"
        assert capi.lz4_decompress(blk, len(data)) == data
        theirs = LZ4.compress(data, asbytes=True) if data else b"\x00"
        assert capi.lz4_decompress(theirs, len(data)) == data
    assert len(capi.lz4_compress(b"\x00" * 100_000)) < 500  # it does compress


def test_malformed_blocks_are_rejected(capi):
    from sph_b200.capi import SphbError

    good = capi.lz4_compress(np.arange(1000, dtype=np.int64).tobytes())
    with pytest.raises(SphbError):
        capi.lz4_decompress(good, 7999)          # output too small
    with pytest.raises(SphbError):
        capi.lz4_decompress(good[:-3], 8000)     # truncated
    with pytest.raises(SphbError):
        capi.lz4_decompress(b"\x0f\x01\x00", 64)  # a match that points in front of the output


def ref_vec_file(vec: np.ndarray, chunks: int = 1) -> bytes:
    """writeCompressedVecToBinary's layout with liblz4 blocks (FileIO.hpp:165-278)."""
    raw = vec.tobytes()
    out = struct.pack("<QQQ", len(raw), vec.size, chunks)
    if chunks == 1:
        blk = LZ4.compress(raw, asbytes=True)
        return out + struct.pack("<i", len(blk)) + blk
    base, rem = divmod(vec.size, chunks)
    at = 0
    for c in range(chunks):
        ln = base + (1 if c < rem else 0)
        part = vec[at:at + ln].tobytes()
        blk = LZ4.compress(part, asbytes=True)
        out += struct.pack("<ii", len(blk), len(part)) + blk
        at += ln
    return out


@pytest.mark.parametrize("dtype", [np.int64, np.float32, np.uint8])
def test_vec_files_both_directions(capi, tmp_path, dtype):
    rng = np.random.default_rng(3)
    vec = (rng.random(20_001) * 1000).astype(dtype)
    # ours -> parsed by hand, block decoded by liblz4
    p = tmp_path / "v.cache"
    capi.cache_write_vec(p, vec)
    blob = p.read_bytes()
    total, items, chunks = struct.unpack_from("<QQQ", blob, 0)
    (csize,) = struct.unpack_from("<i", blob, 24)
    assert (total, items, chunks) == (vec.nbytes, vec.size, 1) and len(blob) == 28 + csize
    assert LZ4.decompress(blob[28:], decompressed_size=total, asbytes=True) == vec.tobytes()
    assert np.array_equal(capi.cache_read_vec(p, dtype), vec)
    # the reference's writer (liblz4 blocks) -> our reader, single and chunked
    for ch in (1, 4):
        q = tmp_path / f"r{ch}.cache"
        q.write_bytes(ref_vec_file(vec, ch))
        assert np.array_equal(capi.cache_read_vec(q, dtype), vec)
    with pytest.raises(Exception):
        capi.cache_read_vec(p, np.int16)  # element size mismatch


def test_vec_chunked_layout(capi, tmp_path):
    vec = np.arange(10_007, dtype=np.int64)
    capi.cache_chunk_limit(30_000)  # 80 056 bytes -> ceil(2.67) * 2 = 6 chunks
    try:
        p = tmp_path / "c.cache"
        capi.cache_write_vec(p, vec)
    finally:
        capi.cache_chunk_limit(0)
    blob = p.read_bytes()
    total, items, chunks = struct.unpack_from("<QQQ", blob, 0)
    assert (total, items, chunks) == (vec.nbytes, vec.size, 6)
    at, got = 24, []
    for c in range(6):
        csize, osize = struct.unpack_from("<ii", blob, at)
        assert osize == (10_007 // 6 + (1 if c < 10_007 % 6 else 0)) * 8  # divide_into_spans (FileIO.cpp:140-155)
        got.append(LZ4.decompress(blob[at + 8:at + 8 + csize], decompressed_size=osize, asbytes=True))
        at += 8 + csize
    assert at == len(blob) and b"".join(got) == vec.tobytes()
    assert np.array_equal(capi.cache_read_vec(p, np.int64), vec)


def test_graph_files(capi, tmp_path):
    rng = np.random.default_rng(5)
    ids = rng.integers(0, 500, (500, 7)).astype(np.int64)
    dist = np.sort(rng.random((500, 7)).astype(np.float32), axis=1)
    base = str(tmp_path / "case_nns_DataKnn")
    capi.cache_write_graph(base, ids, dist, symmetric=False)
    assert np.fromfile(base + "NNs.cache", np.int64).tolist() == [7] * 500  # raw, FileIO.cpp:606-609
    assert np.fromfile(base + "Symmetric.cache", np.uint8).tolist() == [0]
    rp, i2, d2, sym = capi.cache_read_graph(base)
    assert np.array_equal(rp, np.arange(501) * 7) and np.array_equal(i2, ids.reshape(-1)) and np.array_equal(d2, dist.reshape(-1)) and sym is False
    # ragged + symmetric
    nns = rng.integers(1, 12, 300)
    rowptr = np.concatenate([[0], np.cumsum(nns)]).astype(np.int64)
    rid = rng.integers(0, 300, rowptr[-1]).astype(np.int64)
    rdist = rng.random(rowptr[-1]).astype(np.float32)
    base2 = str(tmp_path / "case_nns_SymKnn")
    capi.cache_write_graph(base2, rid, rdist, rowptr=rowptr, symmetric=True)
    rp, i2, d2, sym = capi.cache_read_graph(base2)
    assert np.array_equal(rp, rowptr) and np.array_equal(i2, rid) and np.array_equal(d2, rdist) and sym is True
    # a graph written the reference's way loads
    base3 = str(tmp_path / "ref_")
    (tmp_path / "ref_Distances.cache").write_bytes(ref_vec_file(rdist))
    (tmp_path / "ref_Indices.cache").write_bytes(ref_vec_file(rid, 2))
    nns.astype(np.int64).tofile(base3 + "NNs.cache")
    np.array([1], np.uint8).tofile(base3 + "Symmetric.cache")
    rp, i2, d2, sym = capi.cache_read_graph(base3)
    assert np.array_equal(rp, rowptr) and np.array_equal(i2, rid) and np.array_equal(d2, rdist) and sym is True


def serial_rows(rowptr, col, val, r0, r1):
    out = struct.pack("<Q", r1 - r0)
    for r in range(r0, r1):
        b, e = rowptr[r], rowptr[r + 1]
        out += struct.pack("<Q", e - b)
        pairs = np.empty(e - b, dtype=[("c", "<u4"), ("v", "<f4")])
        pairs["c"], pairs["v"] = col[b:e], val[b:e]
        out += pairs.tobytes()
    return out


def test_prob_files(capi, tmp_path):
    rng = np.random.default_rng(9)
    nnz_row = rng.integers(0, 40, 700)
    rowptr = np.concatenate([[0], np.cumsum(nnz_row)]).astype(np.int64)
    col = np.concatenate([np.sort(rng.choice(700, k, replace=False)) for k in nnz_row]).astype(np.uint32)
    val = rng.random(rowptr[-1]).astype(np.float32)
    p = tmp_path / "case_lls_probDist1.cache"
    capi.cache_write_prob(p, rowptr, col, val)
    blob = p.read_bytes()
    total, rows, chunks = struct.unpack_from("<QQQ", blob, 0)
    (csize,) = struct.unpack_from("<i", blob, 24)
    assert (total, rows, chunks) == (8 + 700 * 8 + int(rowptr[-1]) * 8, 700, 1)
    assert LZ4.decompress(blob[28:28 + csize], decompressed_size=total, asbytes=True) == serial_rows(rowptr, col, val, 0, 700)
    rp, c2, v2 = capi.cache_read_prob(p)
    assert np.array_equal(rp, rowptr) and np.array_equal(c2, col) and np.array_equal(v2, val)
    # chunked: rows split, every chunk with its own row count (FileIO.cpp:741-797)
    capi.cache_chunk_limit(40_000)
    try:
        q = tmp_path / "chunked.cache"
        capi.cache_write_prob(q, rowptr, col, val)
    finally:
        capi.cache_chunk_limit(0)
    blob = q.read_bytes()
    total2, rows2, chunks2 = struct.unpack_from("<QQQ", blob, 0)
    assert total2 == total and rows2 == 700 and chunks2 == 2 * -(-total // 40_000)
    at, r0 = 24, 0
    for c in range(chunks2):
        csize, osize = struct.unpack_from("<II", blob, at)
        ln = 700 // chunks2 + (1 if c < 700 % chunks2 else 0)
        assert LZ4.decompress(blob[at + 8:at + 8 + csize], decompressed_size=osize, asbytes=True) == serial_rows(rowptr, col, val, r0, r0 + ln)
        at += 8 + csize
        r0 += ln
    assert at == len(blob) and r0 == 700
    rp, c2, v2 = capi.cache_read_prob(q)
    assert np.array_equal(rp, rowptr) and np.array_equal(c2, col) and np.array_equal(v2, val)
    # the reference's writer (liblz4 block) -> our reader
    ser = serial_rows(rowptr, col, val, 0, 700)
    blk = LZ4.compress(ser, asbytes=True)
    r = tmp_path / "ref.cache"
    r.write_bytes(struct.pack("<QQQ", len(ser), 700, 1) + struct.pack("<i", len(blk)) + blk)
    rp, c2, v2 = capi.cache_read_prob(r)
    assert np.array_equal(rp, rowptr) and np.array_equal(c2, col) and np.array_equal(v2, val)


def test_missing_and_truncated_files(capi, tmp_path):
    from sph_b200.capi import SphbError

    with pytest.raises(SphbError):
        capi.cache_read_vec(tmp_path / "nope.cache", np.int64)
    p = tmp_path / "t.cache"
    capi.cache_write_vec(p, np.arange(1000, dtype=np.int64))
    p.write_bytes(p.read_bytes()[:-10])
    with pytest.raises(SphbError):
        capi.cache_read_vec(p, np.int64)


def test_host_mirrors_round_trip(capi, tmp_path):
    from sph_b200 import host

    rng = np.random.default_rng(1)
    g = host.Graph(rng.integers(0, 50, (50, 4)).astype(np.int64), rng.random((50, 4)).astype(np.float32))
    host.writeCompressedGraphToBinary(str(tmp_path / "x_nns_DataKnn"), g)
    r = host.loadCompressedGraphFromBinary(str(tmp_path / "x_nns_DataKnn"))
    assert np.array_equal(r.knnIndices, g.knnIndices.reshape(-1)) and np.array_equal(r.knnDistances, g.knnDistances.reshape(-1)) and not r.symmetric
    assert (r.nns == 4).all()
    m = host.SparseRows(np.array([0, 2, 2, 5], np.int64), np.array([1, 2, 0, 1, 2], np.uint32), np.array([.5, .5, .2, .3, .5], np.float32))
    host.writeCompressedSparseMatHDIToBinary(str(tmp_path / "x_lls_probDist1.cache"), m)
    r = host.loadCompressedSparseMatHDIFromBinary(str(tmp_path / "x_lls_probDist1.cache"))
    assert np.array_equal(r.rowptr, m.rowptr) and np.array_equal(r.col, m.col) and np.array_equal(r.val, m.val)
    labels = rng.integers(0, 3, 50).astype(np.int64)
    host.writeCompressedVecToBinary(str(tmp_path / "cc.cache"), labels)
    assert np.array_equal(host.loadCompressedVecFromBinary(str(tmp_path / "cc.cache"), np.int64), labels)


def test_committed_liblz4_blocks_decode(capi, tmp_path):
    """tests/golden/lz4_blocks_liblz4.json: blocks and one whole cache file written by the real liblz4 (make_external_goldens.py);
    this pin does not need pyarrow at test time."""
    import base64
    import hashlib
    import json
    from pathlib import Path

    g = json.loads((Path(__file__).parent / "golden" / "lz4_blocks_liblz4.json").read_text())
    for name, b in g["blocks"].items():
        data = capi.lz4_decompress(base64.b64decode(b["block_b64"]), b["size"])
        assert len(data) == b["size"] and hashlib.sha256(data).hexdigest() == b["sha256"], name
    p = tmp_path / "golden.cache"
    p.write_bytes(base64.b64decode(g["vec_file"]["file_b64"]))
    assert np.array_equal(capi.cache_read_vec(p, np.int64), (np.arange(2500, dtype=np.int64) * 7) % 1000)
