"""ctypes binding of the C ABI (include/sphb.h) — numpy in, numpy out.

This is the thin Python view of the library used by the tests and bench.py; a
C++ caller binds the same symbols through include/sph_b200/*.hpp.  There is no
CPU fallback: if the shared library is missing or no B200 is visible the calls
raise.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import numpy as np

# SPHB_LIB selects another build of the same library (e.g. the -DSPHB_KNN_PROFILE one)
_LIB_PATH = Path(os.environ.get("SPHB_LIB") or (Path(__file__).resolve().parent / "lib" / "libsphb200.so")).resolve()
_lib = None

METRIC_INNER_PRODUCT = 0
METRIC_L2 = 1
METRIC_COSINE = 2
W_CONSTANT, W_LINEAR, W_NORMAL, W_ONLYLAST, W_FIRST_VISIT = range(5)

i64p = C.POINTER(C.c_int64)
u64p = C.POINTER(C.c_uint64)
u32p = C.POINTER(C.c_uint32)
i32p = C.POINTER(C.c_int32)
f32p = C.POINTER(C.c_float)
f64p = C.POINTER(C.c_double)

# name -> (restype, argtypes); must list every symbol include/sphb.h declares
SIGNATURES = {
    "sphb_init": (C.c_int, [C.c_int]),
    "sphb_shutdown": (None, []),
    "sphb_last_error": (C.c_char_p, []),
    "sphb_version": (C.c_int, []),
    "sphb_launch_count": (C.c_int64, []),
    "sphb_reset_launch_count": (None, []),
    "sphb_stream": (C.c_void_p, []),
    "sphb_knn_exact": (C.c_int, [f32p, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int, i64p, f32p]),
    "sphb_knn_exact_slab": (C.c_int, [f32p, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int64, C.c_int64, i64p, f32p]),
    "sphb_knn_exact_dev": (C.c_int, [C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]),
    "sphb_renorm_l2": (C.c_int, [f32p, C.c_int64, C.c_int, f32p]),
    "sphb_rw_post_merge": (C.c_int, [C.c_int64, i64p, u32p, f32p, C.c_int, C.c_int, i64p, u32p, f32p, i64p]),
    "sphb_rw_post_merge_dev": (C.c_int, [C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, i64p,
                                         C.c_void_p]),
    "sphb_knn_last_stats": (C.c_int, [i64p]),
    "sphb_knn_last_timing": (C.c_int, [f32p]),
    "sphb_knn_rows_noalloc_dev": (C.c_int, [C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p,
                                            C.c_void_p]),
    "sphb_knn_last_overflow_rows": (C.c_int, [C.c_int, i64p, C.c_int64, i64p]),
    "sphb_knn_graph_prow": (C.c_int, [f32p, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int64, C.c_int64, i64p, f32p, u32p, f32p]),
    "sphb_knn_graph_prow_multi": (C.c_int, [f32p, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.POINTER(C.c_int), C.c_int, i64p, f32p, u32p,
                                            f32p]),
    "sphb_prow_gaussian_graph": (C.c_int, [i64p, f32p, C.c_int64, C.c_int, C.c_float, u32p, f32p, f64p]),
    "sphb_prow_gaussian_graph_dev": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "sphb_prow_linear_graph": (C.c_int, [i64p, f32p, C.c_int64, C.c_int, u32p, f32p]),
    "sphb_prow_gaussian_inplace": (C.c_int, [C.c_int64, i64p, f32p, C.c_float, C.c_int]),
    "sphb_prow_gaussian_flat": (C.c_int, [f32p, C.c_int64, C.c_int, C.c_double, f32p]),
    "sphb_component_centers": (C.c_int, [f32p, C.c_int64, C.c_int, i64p, C.c_int64, f32p]),
    "sphb_component_centers_dev": (C.c_int, [C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]),
    "sphb_bridge_components": (C.c_int, [f32p, C.c_int64, C.c_int, C.c_int, i64p, C.c_int64, i64p, C.c_int64, i64p, i64p, f32p]),
    "sphb_bridge_components_dev": (C.c_int, [C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_void_p, C.c_int64, i64p, C.c_int64, C.c_void_p,
                                             C.c_void_p, C.c_void_p, C.c_void_p]),
    "sphb_level_knn_euclid": (C.c_int, [f32p, C.c_int64, C.c_int, C.c_int64, i64p, i64p, C.c_int64, i64p, f32p]),
    "sphb_prow_exponential_graph": (C.c_int, [i64p, f32p, C.c_int64, C.c_int, C.c_int, u32p, f32p]),
    "sphb_prow_exponential_graph_dev": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "sphb_prow_exponential_inplace": (C.c_int, [C.c_int64, i64p, u32p, f32p, C.c_int]),
    "sphb_prow_exponential_inplace_dev": (C.c_int, [C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "sphb_symmetrize": (C.c_int, [C.c_int64, i64p, u32p, f32p, C.c_int, i64p, u32p, f32p, C.c_int64, i64p]),
    "sphb_random_walks": (C.c_int, [C.c_int64, i64p, u32p, f32p, C.c_uint64, C.c_uint64, C.c_int, C.c_float, C.c_uint64, C.c_uint64,
                                    C.c_int, C.c_int, i64p, u32p, f32p, C.c_int64, i64p]),
    "sphb_random_walks_slab": (C.c_int, [C.c_int64, i64p, u32p, f32p, C.c_uint64, C.c_uint64, C.c_int, C.c_float, C.c_uint64, C.c_uint64,
                                         C.c_int, C.c_int, C.c_int64, C.c_int64, i64p, u32p, f32p, C.c_int64, i64p]),
    "sphb_rw_merge": (C.c_int, [C.c_int64, i64p, u32p, f32p, u64p, C.c_int64, C.c_int, C.c_int, i64p, u32p, f32p, C.c_int64, i64p]),
    "sphb_create_similarities": (C.c_int, [C.c_int64, i64p, u32p, f32p, C.c_int64, C.c_float, C.c_int, i64p, u32p, f32p, i32p]),
    "sphb_pair_sims": (C.c_int, [C.c_int64, i64p, u32p, f32p, i64p, i64p, C.c_int64, C.c_int, f32p]),
    "sphb_find_k": (C.c_int, [C.c_int64, i64p, u32p, f32p, C.c_int64, C.c_int, u32p, f32p, i32p]),
    "sphb_neighbor_overlap": (C.c_int, [i64p, C.c_int64, C.c_int, C.c_int64, i64p, i64p, i64p, i64p, C.c_int64, f32p]),
    "sphb_symmetrize_graph": (C.c_int, [C.c_int64, i64p, C.c_int, i64p, f32p, i64p, i64p, f32p, C.c_int64, i64p]),
    "sphb_prow_graph_ragged": (C.c_int, [C.c_int64, i64p, i64p, f32p, C.c_float, C.c_int, i64p, u32p, f32p]),
    "sphb_weak_components": (C.c_int, [C.c_int64, i64p, C.c_int, i64p, i64p, i64p]),
    "sphb_symmetrize_graph_dev": (C.c_int, [C.c_int64, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p,
                                            C.c_int64, i64p, C.c_void_p]),
    "sphb_merge_graph_nodes": (C.c_int, [C.c_int64, i64p, C.c_int, i64p, f32p, u64p, C.c_int64, i64p, i64p, f32p, C.c_int64, i64p]),
    "sphb_merge_data_distances": (C.c_int, [C.c_int64, i64p, u32p, f32p, u64p, C.c_int64, i64p, u32p, f32p, C.c_int64, i64p]),
    "sphb_rw_as_knn_distances": (C.c_int, [C.c_int64, i64p, u32p, f32p, i64p, i64p, f32p, i64p]),
    "sphb_remove_zeros": (C.c_int, [C.c_int64, i64p, u32p, f32p, i64p, u32p, f32p, i64p]),
    "sphb_level_knn_walks": (C.c_int, [C.c_int64, i64p, u32p, f32p, C.c_int, C.c_int64, i64p, f32p]),
    "sphb_level_knn_overlap": (C.c_int, [i64p, C.c_int64, C.c_int, C.c_int64, i64p, i64p, C.c_int64, i64p, f32p]),
    "sphb_euclid_hausdorff": (C.c_int, [f32p, C.c_int64, C.c_int, C.c_int64, i64p, i64p, i64p, i64p, C.c_int64, f32p]),
    "sphb_random_walks_dev": (C.c_int, [C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, C.c_int, C.c_float, C.c_uint64, C.c_uint64,
                                        C.c_int, C.c_int, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, i64p, C.c_void_p]),
    "sphb_rw_merge_dev": (C.c_int, [C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_void_p,
                                    C.c_void_p, C.c_void_p, C.c_int64, i64p, C.c_void_p]),
    "sphb_create_similarities_dev": (C.c_int, [C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_float, C.c_int, C.c_void_p,
                                               C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "sphb_prow_gaussian_inplace_dev": (C.c_int, [C.c_int64, C.c_void_p, C.c_void_p, C.c_int, C.c_float, C.c_int, C.c_void_p]),
    "sphb_symmetrize_dev": (C.c_int, [C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                      i64p, C.c_void_p]),
    "sphb_find_k_dev": (C.c_int, [C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int64, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                  C.c_void_p]),
    "sphb_level_prob_dist_dev": (C.c_int, [C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_float, C.c_int, C.c_void_p, C.c_int,
                                           C.c_float, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, i64p, C.c_void_p]),
    "sphb_level_prob_dist": (C.c_int, [C.c_int64, i64p, u32p, f32p, C.c_int64, C.c_float, C.c_int, i64p, C.c_int, C.c_float, C.c_int, C.c_int, i64p,
                                       u32p, f32p, C.c_int64, i64p]),
    "sphb_level_last_timing": (C.c_int, [f32p]),
    "sphb_create_similarities_slab": (C.c_int, [C.c_int64, i64p, u32p, f32p, C.c_int64, C.c_float, C.c_int, i64p, C.c_int64, C.c_int64, u32p, f32p, i32p]),
    "sphb_create_similarities_slab_dev": (C.c_int, [C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_float, C.c_int, C.c_void_p,
                                                    C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "sphb_level_from_sims_dev": (C.c_int, [C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                           C.c_void_p, C.c_int64, i64p, C.c_void_p]),
    "sphb_geodesic_distances": (C.c_int, [C.c_int64, i64p, C.c_int, i64p, f32p, i64p, i64p, i64p, C.c_int64, f32p]),
    "sphb_geodesic_hausdorff": (C.c_int, [C.c_int64, i64p, C.c_int, i64p, f32p, i64p, C.c_int64, i64p, i64p, i64p, i64p, C.c_int64, f32p]),
    "sphb_geodesic_last_stats": (C.c_int, [i64p]),
    "sphb_geodesic_graph_create": (C.c_int, [C.c_int64, i64p, C.c_int, i64p, f32p, i64p, C.POINTER(C.c_void_p)]),
    "sphb_geodesic_graph_destroy": (C.c_int, [C.c_void_p]),
    "sphb_geodesic_distances_g": (C.c_int, [C.c_void_p, i64p, i64p, C.c_int64, f32p]),
    "sphb_geodesic_hausdorff_g": (C.c_int, [C.c_void_p, C.c_int64, i64p, i64p, i64p, i64p, C.c_int64, f32p]),
    "sphb_lz4_compress_bound": (C.c_int64, [C.c_int64]),
    "sphb_lz4_compress": (C.c_int64, [C.c_void_p, C.c_int64, C.c_void_p, C.c_int64]),
    "sphb_lz4_decompress": (C.c_int64, [C.c_void_p, C.c_int64, C.c_void_p, C.c_int64]),
    "sphb_cache_write_vec": (C.c_int, [C.c_char_p, C.c_void_p, C.c_int64, C.c_int]),
    "sphb_cache_read_vec": (C.c_int, [C.c_char_p, C.c_void_p, C.c_int64, C.c_int, i64p]),
    "sphb_cache_write_graph": (C.c_int, [C.c_char_p, C.c_int64, i64p, C.c_int, i64p, f32p, C.c_int]),
    "sphb_cache_read_graph": (C.c_int, [C.c_char_p, C.c_int64, C.c_int64, i64p, i64p, i64p, i64p, f32p, C.POINTER(C.c_int)]),
    "sphb_cache_write_prob": (C.c_int, [C.c_char_p, C.c_int64, i64p, u32p, f32p]),
    "sphb_cache_read_prob": (C.c_int, [C.c_char_p, C.c_int64, C.c_int64, i64p, i64p, i64p, u32p, f32p]),
}


class SphbError(RuntimeError):
    pass


def lib_path() -> Path:
    return _LIB_PATH


def load():
    """Loads the C-ABI library (no compute, no GPU needed)."""
    global _lib
    if _lib is not None:
        return _lib
    if not _LIB_PATH.exists():
        raise SphbError(f"{_LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` (there is no CPU fallback)")
    lib = C.CDLL(str(_LIB_PATH))
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    # testing hook, not part of the public header
    lib.sphb_debug_force_knn_path.restype = None
    lib.sphb_debug_force_knn_path.argtypes = [C.c_int]
    lib.sphb_debug_force_sims_slab.restype = None
    lib.sphb_debug_force_sims_slab.argtypes = [C.c_int]
    lib.sphb_debug_force_walk_scan.restype = None
    lib.sphb_debug_force_walk_scan.argtypes = [C.c_int]
    lib.sphb_debug_cache_chunk_limit.restype = None
    lib.sphb_debug_cache_chunk_limit.argtypes = [C.c_int64]
    _lib = lib
    return lib


def _check(rc: int):
    if rc != 0:
        raise SphbError(f"sphb error {rc}: {load().sphb_last_error().decode()}")


def _p(a: np.ndarray, t):
    return a.ctypes.data_as(t)


def init(device: int = -1):
    _check(load().sphb_init(device))


def launch_count() -> int:
    return int(load().sphb_launch_count())


def reset_launch_count():
    load().sphb_reset_launch_count()


def stream_ptr() -> int:
    return int(load().sphb_stream() or 0)


def force_knn_path(path: int):
    """0 auto, 1 SIMT exact kernel, 2 tcgen05 (auto), 3 streaming 3-pass, 4 resident MH=1 3-pass, 5 3-pass (auto flavour), 6 1-pass, 7 1-pass streaming (tests only)."""
    load().sphb_debug_force_knn_path(path)


def force_sims_slab(on: bool):
    """createSimilarities: run the L2-slab accumulator kernel even where the shared-memory row fits (tests only)."""
    load().sphb_debug_force_sims_slab(int(on))


def force_walk_scan(on: bool):
    """doRandomWalks: run the serial-scan kernel instead of the binary-search kernel (tests only)."""
    load().sphb_debug_force_walk_scan(int(on))


def knn_last_stats() -> dict:
    out = np.zeros(8, np.int64)
    _check(load().sphb_knn_last_stats(_p(out, i64p)))
    return dict(tensor_core=int(out[0]), compactions=int(out[1]), accepted=int(out[2]), kp=int(out[3]), overflow_rows=int(out[4]), q_tiles=int(out[5]),
                passes=int(out[6]), resweep_rows=int(out[7]))


def knn_last_overflow_rows(which: int = 0, cap: int = 1 << 20) -> np.ndarray:
    """Global ids of the rows of the last kNN call that were re-swept with the 3-product filter (which = 0) or recomputed by
    the exact SIMT kernel (which = 1)."""
    cnt = np.zeros(1, np.int64)
    out = np.zeros(max(1, cap), np.int64)
    _check(load().sphb_knn_last_overflow_rows(which, _p(out, i64p), cap, _p(cnt, i64p)))
    return out[: min(cap, int(cnt[0]))].copy()


def knn_last_timing() -> dict:
    out = np.zeros(4, np.float32)
    _check(load().sphb_knn_last_timing(_p(out, f32p)))
    return dict(prep_ms=float(out[0]), sweep_ms=float(out[1]), rerank_ms=float(out[2]), fallback_ms=float(out[3]))


def knn_graph_prow(x: np.ndarray, k: int, perplexity: float = 0.0, metric: int = METRIC_L2, l2_squared: bool = False, q0: int = 0,
                   q1: int | None = None, out=None):
    """Stage A + level-0 P rows through the host-buffer entry point; `out` = (ids, dist, cols, vals) to reuse (e.g. pinned) buffers."""
    n, d = x.shape
    q1 = n if q1 is None else q1
    nq = q1 - q0
    if out is None:
        out = (np.empty((nq, k), np.int64), np.empty((nq, k), np.float32), np.empty((nq, k - 1), np.uint32), np.empty((nq, k - 1), np.float32))
    ids, dist, cols, vals = out
    _check(load().sphb_knn_graph_prow(_p(x, f32p), n, d, k, metric, int(l2_squared), perplexity, q0, q1, _p(ids, i64p), _p(dist, f32p),
                                      _p(cols, u32p), _p(vals, f32p)))
    return out


# ------------------------------------------------------------------ Stage A
def knn_graph_prow_multi(x: np.ndarray, k: int, devices, perplexity: float = 0.0, metric: int = METRIC_L2, l2_squared: bool = False, out=None):
    """Stage A + level-0 P rows of the whole stack on the listed GPUs from one call (one worker thread per device)."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    n, d = x.shape
    if out is None:
        out = (np.full((n, k), -1, np.int64), np.full((n, k), -1.0, np.float32), np.zeros((n, k - 1), np.uint32), np.zeros((n, k - 1), np.float32))
    ids, dist, cols, vals = out
    dev = (C.c_int * len(devices))(*devices)
    _check(load().sphb_knn_graph_prow_multi(_p(x, f32p), n, d, k, metric, int(l2_squared), perplexity, dev, len(devices), _p(ids, i64p), _p(dist, f32p),
                                            _p(cols, u32p), _p(vals, f32p)))
    return ids, dist, cols, vals


def knn_exact(x: np.ndarray, k: int, metric: int = METRIC_L2, l2_squared: bool = False, q0: int = 0, q1: int | None = None):
    """Exact kNN of rows [q0,q1) of x against all rows; returns (ids int64 [nq,k], dist float32 [nq,k])."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    n, d = x.shape
    q1 = n if q1 is None else q1
    nq = q1 - q0
    ids = np.full((nq, k), -1, np.int64)
    dist = np.full((nq, k), -1.0, np.float32)
    _check(load().sphb_knn_exact_slab(_p(x, f32p), n, d, k, metric, int(l2_squared), q0, q1, _p(ids, i64p), _p(dist, f32p)))
    return ids, dist


def knn_exact_dev(x_ptr: int, n: int, d: int, k: int, metric: int, l2_squared: bool, q0: int, q1: int, ids_ptr: int, dist_ptr: int, stream: int):
    _check(load().sphb_knn_exact_dev(x_ptr, n, d, k, metric, int(l2_squared), q0, q1, ids_ptr, dist_ptr, stream))


# ------------------------------------------------------------------ Stage B
def prow_gaussian_graph(ids: np.ndarray, dist: np.ndarray, perplexity: float = 0.0, want_sigma: bool = False):
    ids = np.ascontiguousarray(ids, np.int64)
    dist = np.ascontiguousarray(dist, np.float32)
    n, k = ids.shape
    cols = np.zeros((n, k - 1), np.uint32)
    vals = np.zeros((n, k - 1), np.float32)
    sigma = np.zeros(n, np.float64) if want_sigma else None
    _check(load().sphb_prow_gaussian_graph(_p(ids, i64p), _p(dist, f32p), n, k, perplexity, _p(cols, u32p), _p(vals, f32p),
                                           _p(sigma, f64p) if want_sigma else None))
    return (cols, vals, sigma) if want_sigma else (cols, vals)


def prow_gaussian_graph_dev(ids_ptr: int, dist_ptr: int, n: int, k: int, perplexity: float, cols_ptr: int, vals_ptr: int, stream: int):
    _check(load().sphb_prow_gaussian_graph_dev(ids_ptr, dist_ptr, n, k, perplexity, cols_ptr, vals_ptr, None, stream))


def prow_linear_graph(ids: np.ndarray, dist: np.ndarray):
    ids = np.ascontiguousarray(ids, np.int64)
    dist = np.ascontiguousarray(dist, np.float32)
    n, k = ids.shape
    cols = np.zeros((n, k - 1), np.uint32)
    vals = np.zeros((n, k - 1), np.float32)
    _check(load().sphb_prow_linear_graph(_p(ids, i64p), _p(dist, f32p), n, k, _p(cols, u32p), _p(vals, f32p)))
    return cols, vals


def prow_gaussian_inplace(rowptr: np.ndarray, val: np.ndarray, perplexity: float = 0.0, ignore: int = -1) -> np.ndarray:
    rowptr = np.ascontiguousarray(rowptr, np.int64)
    out = np.array(val, dtype=np.float32, copy=True)
    _check(load().sphb_prow_gaussian_inplace(len(rowptr) - 1, _p(rowptr, i64p), _p(out, f32p), perplexity, ignore))
    return out


def prow_exponential_graph(ids: np.ndarray, dist: np.ndarray, sort_cols: bool = False):
    """utils::computeExponentialDistributions(graph, rows): UMAP membership rows of neighbours 1..k-1 -> (cols, vals) [n, k-1]."""
    ids = np.ascontiguousarray(ids, np.int64)
    dist = np.ascontiguousarray(dist, np.float32)
    n, k = ids.shape
    cols = np.zeros((n, k - 1), np.uint32)
    vals = np.zeros((n, k - 1), np.float32)
    _check(load().sphb_prow_exponential_graph(_p(ids, i64p), _p(dist, f32p), n, k, int(sort_cols), _p(cols, u32p), _p(vals, f32p)))
    return cols, vals


def prow_exponential_graph_dev(ids_ptr: int, dist_ptr: int, n: int, k: int, sort_cols: bool, cols_ptr: int, vals_ptr: int, stream: int):
    _check(load().sphb_prow_exponential_graph_dev(ids_ptr, dist_ptr, n, k, int(sort_cols), cols_ptr, vals_ptr, stream))


def prow_exponential_inplace(rowptr: np.ndarray, val: np.ndarray, col: np.ndarray | None = None, ignore: int = -1) -> np.ndarray:
    """utils::computeExponentialDistributions(rows, ignore) on CSR rows; returns the new values."""
    rowptr = np.ascontiguousarray(rowptr, np.int64)
    out = np.array(val, np.float32, copy=True)
    colp = _p(np.ascontiguousarray(col, np.uint32), u32p) if col is not None else None
    _check(load().sphb_prow_exponential_inplace(len(rowptr) - 1, _p(rowptr, i64p), colp, _p(out, f32p), ignore))
    return out


def prow_exponential_inplace_dev(n, rowptr_ptr, col_ptr, val_ptr, max_row_nnz, ignore, stream):
    _check(load().sphb_prow_exponential_inplace_dev(n, rowptr_ptr, col_ptr, val_ptr, max_row_nnz, ignore, stream))


def prow_gaussian_flat(dist: np.ndarray, perplexity: float) -> np.ndarray:
    dist = np.ascontiguousarray(dist, np.float32)
    n, k = dist.shape
    temp = np.zeros((n, k), np.float32)
    _check(load().sphb_prow_gaussian_flat(_p(dist, f32p), n, k, float(perplexity), _p(temp, f32p)))
    return temp


def symmetrize(rowptr, col, val, umap: bool = False):
    rowptr = np.ascontiguousarray(rowptr, np.int64)
    col = np.ascontiguousarray(col, np.uint32)
    val = np.ascontiguousarray(val, np.float32)
    n = len(rowptr) - 1
    cap = max(1, 2 * len(col))
    orp = np.zeros(n + 1, np.int64)
    oc = np.zeros(cap, np.uint32)
    ov = np.zeros(cap, np.float32)
    nnz = np.zeros(1, np.int64)
    _check(load().sphb_symmetrize(n, _p(rowptr, i64p), _p(col, u32p), _p(val, f32p), int(umap), _p(orp, i64p), _p(oc, u32p), _p(ov, f32p), cap, _p(nnz, i64p)))
    m = int(nnz[0])
    return orp, oc[:m].copy(), ov[:m].copy()


# ------------------------------------------------------------------ Stage C
def random_walks(rowptr, col, val, num_walks: int, walk_len: int, weighting: int = W_CONSTANT, prune_value: float = 0.0, prune_steps: int = 0,
                 seed: int = 0, remove_diagonal: bool = True, normalize: bool = True, cap: int | None = None):
    rowptr = np.ascontiguousarray(rowptr, np.int64)
    col = np.ascontiguousarray(col, np.uint32)
    val = np.ascontiguousarray(val, np.float32)
    n = len(rowptr) - 1
    cap = max(1, n * min(n, num_walks * walk_len)) if cap is None else int(cap)
    orp = np.zeros(n + 1, np.int64)
    oc = np.zeros(cap, np.uint32)
    ov = np.zeros(cap, np.float32)
    nnz = np.zeros(1, np.int64)
    _check(load().sphb_random_walks(n, _p(rowptr, i64p), _p(col, u32p), _p(val, f32p), num_walks, walk_len, weighting, prune_value, prune_steps,
                                    seed, int(remove_diagonal), int(normalize), _p(orp, i64p), _p(oc, u32p), _p(ov, f32p), cap, _p(nnz, i64p)))
    m = int(nnz[0])
    return orp, oc[:m].copy(), ov[:m].copy()


def random_walks_slab(rowptr, col, val, num_walks: int, walk_len: int, q0: int, q1: int, weighting: int = W_CONSTANT, prune_value: float = 0.0,
                      prune_steps: int = 0, seed: int = 0, remove_diagonal: bool = True, normalize: bool = True, cap: int | None = None):
    """Random walks of the start points [q0, q1) (multi-GPU shard): CSR of q1 - q0 rows."""
    rowptr = np.ascontiguousarray(rowptr, np.int64)
    col = np.ascontiguousarray(col, np.uint32)
    val = np.ascontiguousarray(val, np.float32)
    n = len(rowptr) - 1
    nq = q1 - q0
    cap = max(1, nq * min(n, num_walks * walk_len)) if cap is None else int(cap)
    orp = np.zeros(nq + 1, np.int64)
    oc = np.zeros(cap, np.uint32)
    ov = np.zeros(cap, np.float32)
    nnz = np.zeros(1, np.int64)
    _check(load().sphb_random_walks_slab(n, _p(rowptr, i64p), _p(col, u32p), _p(val, f32p), num_walks, walk_len, weighting, prune_value, prune_steps,
                                         seed, int(remove_diagonal), int(normalize), q0, q1, _p(orp, i64p), _p(oc, u32p), _p(ov, f32p), cap, _p(nnz, i64p)))
    m = int(nnz[0])
    return orp, oc[:m].copy(), ov[:m].copy()


def rw_merge(rowptr, col, val, parents, n_merged: int, norm: bool = True, weight_by_size: bool = True):
    rowptr = np.ascontiguousarray(rowptr, np.int64)
    col = np.ascontiguousarray(col, np.uint32)
    val = np.ascontiguousarray(val, np.float32)
    parents = np.ascontiguousarray(parents, np.uint64)
    n = len(rowptr) - 1
    cap = max(1, len(col))
    orp = np.zeros(n_merged + 1, np.int64)
    oc = np.zeros(cap, np.uint32)
    ov = np.zeros(cap, np.float32)
    nnz = np.zeros(1, np.int64)
    _check(load().sphb_rw_merge(n, _p(rowptr, i64p), _p(col, u32p), _p(val, f32p), _p(parents, u64p), n_merged, int(norm), int(weight_by_size),
                                _p(orp, i64p), _p(oc, u32p), _p(ov, f32p), cap, _p(nnz, i64p)))
    m = int(nnz[0])
    return orp, oc[:m].copy(), ov[:m].copy()


# ------------------------------------------------------------------ Stage D
def create_similarities(rowptr, col, val, k: int, prune_val: float = 1e-4, variant: int = 0, comp_sizes=None):
    rowptr = np.ascontiguousarray(rowptr, np.int64)
    col = np.ascontiguousarray(col, np.uint32)
    val = np.ascontiguousarray(val, np.float32)
    n = len(rowptr) - 1
    oc = np.zeros((n, k), np.uint32)
    ov = np.zeros((n, k), np.float32)
    nnz = np.zeros(n, np.int32)
    cs = None if comp_sizes is None else np.ascontiguousarray(comp_sizes, np.int64)
    _check(load().sphb_create_similarities(n, _p(rowptr, i64p), _p(col, u32p), _p(val, f32p), k, prune_val, variant,
                                           _p(cs, i64p) if cs is not None else None, _p(oc, u32p), _p(ov, f32p), _p(nnz, i32p)))
    return oc, ov, nnz


def create_similarities_slab(rowptr, col, val, k: int, r0: int, r1: int, prune_val: float = 1e-4, variant: int = 0, comp_sizes=None):
    """Rows [r0, r1) of createSimilarities (D1 sharded by output row); outputs hold r1 - r0 rows."""
    rowptr = np.ascontiguousarray(rowptr, np.int64)
    col = np.ascontiguousarray(col, np.uint32)
    val = np.ascontiguousarray(val, np.float32)
    n = len(rowptr) - 1
    nr = max(r1 - r0, 0)
    oc = np.zeros((nr, k), np.uint32)
    ov = np.zeros((nr, k), np.float32)
    nnz = np.zeros(max(nr, 1), np.int32)
    cs = None if comp_sizes is None else np.ascontiguousarray(comp_sizes, np.int64)
    _check(load().sphb_create_similarities_slab(n, _p(rowptr, i64p), _p(col, u32p), _p(val, f32p), k, prune_val, variant,
                                                _p(cs, i64p) if cs is not None else None, r0, r1, _p(oc, u32p), _p(ov, f32p), _p(nnz, i32p)))
    return oc, ov, nnz[:nr]


def pair_sims(rowptr, col, val, pa, pb, mode: int = 0) -> np.ndarray:
    rowptr = np.ascontiguousarray(rowptr, np.int64)
    col = np.ascontiguousarray(col, np.uint32)
    val = np.ascontiguousarray(val, np.float32)
    pa = np.ascontiguousarray(pa, np.int64)
    pb = np.ascontiguousarray(pb, np.int64)
    out = np.zeros(len(pa), np.float32)
    _check(load().sphb_pair_sims(len(rowptr) - 1, _p(rowptr, i64p), _p(col, u32p), _p(val, f32p), _p(pa, i64p), _p(pb, i64p), len(pa), mode, _p(out, f32p)))
    return out


def find_k(rowptr, col, val, k: int, top: bool):
    """Batched findTopK / findBottomK: (cols [n,k] padded 0xFFFFFFFF, vals [n,k], nnz [n])."""
    rowptr = np.ascontiguousarray(rowptr, np.int64)
    col = np.ascontiguousarray(col, np.uint32)
    val = np.ascontiguousarray(val, np.float32)
    n = len(rowptr) - 1
    oc = np.zeros((n, k), np.uint32)
    ov = np.zeros((n, k), np.float32)
    nnz = np.zeros(n, np.int32)
    _check(load().sphb_find_k(n, _p(rowptr, i64p), _p(col, u32p), _p(val, f32p), k, int(top), _p(oc, u32p), _p(ov, f32p), _p(nnz, i32p)))
    return oc, ov, nnz


def neighbor_overlap(knn_ids, comp_ptr, comp_pix, pa, pb) -> np.ndarray:
    knn_ids = np.ascontiguousarray(knn_ids, np.int64)
    comp_ptr = np.ascontiguousarray(comp_ptr, np.int64)
    comp_pix = np.ascontiguousarray(comp_pix, np.int64)
    pa = np.ascontiguousarray(pa, np.int64)
    pb = np.ascontiguousarray(pb, np.int64)
    out = np.zeros(len(pa), np.float32)
    n, k = knn_ids.shape
    _check(load().sphb_neighbor_overlap(_p(knn_ids, i64p), n, k, len(comp_ptr) - 1, _p(comp_ptr, i64p), _p(comp_pix, i64p), _p(pa, i64p),
                                        _p(pb, i64p), len(pa), _p(out, f32p)))
    return out


# ------------------------------------------------------------------ ragged-graph passes
def _graph_args(rowptr, ids, dist):
    """(n, rowptr or None, k, ids, dist): a 2-D ids array is a fixed-k graph, otherwise rowptr gives the rows."""
    ids = np.ascontiguousarray(ids, np.int64)
    dist = np.ascontiguousarray(dist, np.float32)
    if rowptr is None:
        n, k = ids.shape
        return n, None, k, ids, dist
    rowptr = np.ascontiguousarray(rowptr, np.int64)
    return len(rowptr) - 1, rowptr, 0, ids, dist


def level_knn_euclid(x: np.ndarray, comp_ptr, comp_pix, k: int):
    """Exact level kNN in the Euclid (Hausdorff) component metric: (ids int64 [ncomp, k], dist float32 [ncomp, k])."""
    x = np.ascontiguousarray(x, np.float32)
    comp_ptr = np.ascontiguousarray(comp_ptr, np.int64)
    comp_pix = np.ascontiguousarray(comp_pix, np.int64)
    n, d = x.shape
    ncomp = len(comp_ptr) - 1
    ids = np.full((ncomp, k), -1, np.int64)
    dist = np.full((ncomp, k), -1.0, np.float32)
    _check(load().sphb_level_knn_euclid(_p(x, f32p), n, d, ncomp, _p(comp_ptr, i64p), _p(comp_pix, i64p), k, _p(ids, i64p), _p(dist, f32p)))
    return ids, dist


def component_centers(x: np.ndarray, labels: np.ndarray, num_components: int) -> np.ndarray:
    """Centre of mass of every connected component (connectComponents' computeCentersOfMass): [num_components, d] float32."""
    x = np.ascontiguousarray(x, np.float32)
    labels = np.ascontiguousarray(labels, np.int64)
    n, d = x.shape
    centers = np.zeros((num_components, d), np.float32)
    _check(load().sphb_component_centers(_p(x, f32p), n, d, _p(labels, i64p), num_components, _p(centers, f32p)))
    return centers


def bridge_components(x: np.ndarray, labels: np.ndarray, num_components: int, edges: np.ndarray, metric: int = METRIC_L2):
    """Closest pair (point of A, point of B, squared L2) for every (component A, component B) edge: (ids_a, ids_b, dist)."""
    x = np.ascontiguousarray(x, np.float32)
    labels = np.ascontiguousarray(labels, np.int64)
    edges = np.ascontiguousarray(edges, np.int64).reshape(-1, 2)
    n, d = x.shape
    m = len(edges)
    ia = np.full(m, -1, np.int64)
    ib = np.full(m, -1, np.int64)
    dist = np.zeros(m, np.float32)
    _check(load().sphb_bridge_components(_p(x, f32p), n, d, metric, _p(labels, i64p), num_components, _p(edges, i64p), m, _p(ia, i64p),
                                         _p(ib, i64p), _p(dist, f32p)))
    return ia, ib, dist


def symmetrize_graph(ids, dist, rowptr=None):
    """utils::symmetrizeGraph: returns (rowptr, ids, dist) of the ragged symmetric graph."""
    n, rp, k, ids, dist = _graph_args(rowptr, ids, dist)
    cap = max(1, 2 * ids.size)
    orp = np.zeros(n + 1, np.int64)
    oi = np.zeros(cap, np.int64)
    od = np.zeros(cap, np.float32)
    nnz = np.zeros(1, np.int64)
    _check(load().sphb_symmetrize_graph(n, _p(rp, i64p) if rp is not None else None, k, _p(ids, i64p), _p(dist, f32p), _p(orp, i64p), _p(oi, i64p),
                                        _p(od, f32p), cap, _p(nnz, i64p)))
    m = int(nnz[0])
    return orp, oi[:m].copy(), od[:m].copy()


def prow_graph_ragged(rowptr, ids, dist, perplexity: float = 0.0, linear: bool = False):
    """Gaussian / linear probability rows of a ragged graph: CSR (rowptr, cols, vals)."""
    rowptr = np.ascontiguousarray(rowptr, np.int64)
    ids = np.ascontiguousarray(ids, np.int64)
    dist = np.ascontiguousarray(dist, np.float32)
    n = len(rowptr) - 1
    m = max(int(rowptr[-1]) - n, 0)
    orp = np.zeros(n + 1, np.int64)
    oc = np.zeros(max(m, 1), np.uint32)
    ov = np.zeros(max(m, 1), np.float32)
    _check(load().sphb_prow_graph_ragged(n, _p(rowptr, i64p), _p(ids, i64p), _p(dist, f32p), perplexity, int(linear), _p(orp, i64p), _p(oc, u32p),
                                         _p(ov, f32p)))
    return orp, oc[:m].copy(), ov[:m].copy()


def weak_components(ids, rowptr=None):
    """Connected components (edges undirected): (num_components, labels int64 [n])."""
    ids = np.ascontiguousarray(ids, np.int64)
    if rowptr is None:
        n, k = ids.shape
        rp = None
    else:
        rp = np.ascontiguousarray(rowptr, np.int64)
        n, k = len(rp) - 1, 0
    labels = np.zeros(max(n, 1), np.int64)
    num = np.zeros(1, np.int64)
    _check(load().sphb_weak_components(n, _p(rp, i64p) if rp is not None else None, k, _p(ids, i64p), _p(labels, i64p), _p(num, i64p)))
    return int(num[0]), labels[:n]


def symmetrize_graph_dev(n: int, rowptr_ptr: int, k: int, ids_ptr: int, dist_ptr: int, nnz: int, out_rowptr_ptr: int, out_ids_ptr: int,
                         out_dist_ptr: int, cap: int, stream: int) -> int:
    """Device-pointer symmetrizeGraph; returns the number of entries written."""
    out = np.zeros(1, np.int64)
    _check(load().sphb_symmetrize_graph_dev(n, rowptr_ptr or None, k, ids_ptr, dist_ptr, nnz, out_rowptr_ptr, out_ids_ptr, out_dist_ptr, cap,
                                            _p(out, i64p), stream or None))
    return int(out[0])


def merge_graph_nodes(ids, dist, parents, n_merged: int, rowptr=None):
    """utils::mergeGraphNodes: returns (rowptr, ids, dist) of the merged graph."""
    n, rp, k, ids, dist = _graph_args(rowptr, ids, dist)
    parents = np.ascontiguousarray(parents, np.uint64)
    cap = max(1, ids.size)
    orp = np.zeros(n_merged + 1, np.int64)
    oi = np.zeros(cap, np.int64)
    od = np.zeros(cap, np.float32)
    nnz = np.zeros(1, np.int64)
    _check(load().sphb_merge_graph_nodes(n, _p(rp, i64p) if rp is not None else None, k, _p(ids, i64p), _p(dist, f32p), _p(parents, u64p), n_merged,
                                         _p(orp, i64p), _p(oi, i64p), _p(od, f32p), cap, _p(nnz, i64p)))
    m = int(nnz[0])
    return orp, oi[:m].copy(), od[:m].copy()


def merge_data_distances(rowptr, col, val, parents, n_merged: int):
    rowptr = np.ascontiguousarray(rowptr, np.int64)
    col = np.ascontiguousarray(col, np.uint32)
    val = np.ascontiguousarray(val, np.float32)
    parents = np.ascontiguousarray(parents, np.uint64)
    n = len(rowptr) - 1
    cap = max(1, len(col))
    orp = np.zeros(n_merged + 1, np.int64)
    oc = np.zeros(cap, np.uint32)
    ov = np.zeros(cap, np.float32)
    nnz = np.zeros(1, np.int64)
    _check(load().sphb_merge_data_distances(n, _p(rowptr, i64p), _p(col, u32p), _p(val, f32p), _p(parents, u64p), n_merged, _p(orp, i64p),
                                            _p(oc, u32p), _p(ov, f32p), cap, _p(nnz, i64p)))
    m = int(nnz[0])
    return orp, oc[:m].copy(), ov[:m].copy()


def rw_as_knn_distances(rowptr, col, val):
    rowptr = np.ascontiguousarray(rowptr, np.int64)
    col = np.ascontiguousarray(col, np.uint32)
    val = np.ascontiguousarray(val, np.float32)
    n = len(rowptr) - 1
    cap = max(1, len(col))
    orp = np.zeros(n + 1, np.int64)
    oi = np.zeros(cap, np.int64)
    od = np.zeros(cap, np.float32)
    nnz = np.zeros(1, np.int64)
    _check(load().sphb_rw_as_knn_distances(n, _p(rowptr, i64p), _p(col, u32p), _p(val, f32p), _p(orp, i64p), _p(oi, i64p), _p(od, f32p), _p(nnz, i64p)))
    m = int(nnz[0])
    return orp, oi[:m].copy(), od[:m].copy()


def remove_zeros(rowptr, col, val):
    rowptr = np.ascontiguousarray(rowptr, np.int64)
    col = np.ascontiguousarray(col, np.uint32)
    val = np.ascontiguousarray(val, np.float32)
    n = len(rowptr) - 1
    cap = max(1, len(col))
    orp = np.zeros(n + 1, np.int64)
    oc = np.zeros(cap, np.uint32)
    ov = np.zeros(cap, np.float32)
    nnz = np.zeros(1, np.int64)
    _check(load().sphb_remove_zeros(n, _p(rowptr, i64p), _p(col, u32p), _p(val, f32p), _p(orp, i64p), _p(oc, u32p), _p(ov, f32p), _p(nnz, i64p)))
    m = int(nnz[0])
    return orp, oc[:m].copy(), ov[:m].copy()


def level_knn_walks(rowptr, col, val, k: int, mode: int = 0):
    """Exact level kNN in the random-walk metric (mode 0 Bhattacharyya, 1 single overlap): (ids [n,k], dist [n,k])."""
    rowptr = np.ascontiguousarray(rowptr, np.int64)
    col = np.ascontiguousarray(col, np.uint32)
    val = np.ascontiguousarray(val, np.float32)
    n = len(rowptr) - 1
    ids = np.full((n, k), -1, np.int64)
    dist = np.full((n, k), -1.0, np.float32)
    _check(load().sphb_level_knn_walks(n, _p(rowptr, i64p), _p(col, u32p), _p(val, f32p), mode, k, _p(ids, i64p), _p(dist, f32p)))
    return ids, dist


def level_knn_overlap(knn_ids, comp_ptr, comp_pix, k: int):
    """Exact level kNN in the neighbour-overlap metric: (ids [ncomp,k], dist [ncomp,k])."""
    knn_ids = np.ascontiguousarray(knn_ids, np.int64)
    comp_ptr = np.ascontiguousarray(comp_ptr, np.int64)
    comp_pix = np.ascontiguousarray(comp_pix, np.int64)
    n, kd = knn_ids.shape
    ncomp = len(comp_ptr) - 1
    ids = np.full((ncomp, k), -1, np.int64)
    dist = np.full((ncomp, k), -1.0, np.float32)
    _check(load().sphb_level_knn_overlap(_p(knn_ids, i64p), n, kd, ncomp, _p(comp_ptr, i64p), _p(comp_pix, i64p), k, _p(ids, i64p), _p(dist, f32p)))
    return ids, dist


def euclid_hausdorff(x, comp_ptr, comp_pix, pa, pb) -> np.ndarray:
    x = np.ascontiguousarray(x, np.float32)
    comp_ptr = np.ascontiguousarray(comp_ptr, np.int64)
    comp_pix = np.ascontiguousarray(comp_pix, np.int64)
    pa = np.ascontiguousarray(pa, np.int64)
    pb = np.ascontiguousarray(pb, np.int64)
    out = np.zeros(len(pa), np.float32)
    n, d = x.shape
    _check(load().sphb_euclid_hausdorff(_p(x, f32p), n, d, len(comp_ptr) - 1, _p(comp_ptr, i64p), _p(comp_pix, i64p), _p(pa, i64p), _p(pb, i64p),
                                        len(pa), _p(out, f32p)))
    return out


def _geo_graph(ids, dist, rowptr):
    ids = np.ascontiguousarray(ids, np.int64)
    dist = np.ascontiguousarray(dist, np.float32)
    if rowptr is None:
        n, k = ids.shape
        return n, None, k, ids, dist
    rp = np.ascontiguousarray(rowptr, np.int64)
    return len(rp) - 1, rp, 0, ids, dist


def geodesic_distances(ids, dist, start, end, rowptr=None, labels=None) -> np.ndarray:
    """computeShortestPath (sph/utils/ShortestPath.cpp:100-171) for point pairs; -1 where there is no path."""
    n, rp, k, ids, dist = _geo_graph(ids, dist, rowptr)
    start = np.ascontiguousarray(start, np.int64)
    end = np.ascontiguousarray(end, np.int64)
    lab = None if labels is None else np.ascontiguousarray(labels, np.int64)
    out = np.zeros(len(start), np.float32)
    _check(load().sphb_geodesic_distances(n, None if rp is None else _p(rp, i64p), k, _p(ids, i64p), _p(dist, f32p), None if lab is None else _p(lab, i64p),
                                          _p(start, i64p), _p(end, i64p), len(start), _p(out, f32p)))
    return out


def geodesic_hausdorff(ids, dist, comp_ptr, comp_pix, pa, pb, rowptr=None, labels=None) -> np.ndarray:
    """geodesicDistance (sph/utils/Similarities.cpp:234-309) for component pairs."""
    n, rp, k, ids, dist = _geo_graph(ids, dist, rowptr)
    comp_ptr = np.ascontiguousarray(comp_ptr, np.int64)
    comp_pix = np.ascontiguousarray(comp_pix, np.int64)
    pa = np.ascontiguousarray(pa, np.int64)
    pb = np.ascontiguousarray(pb, np.int64)
    lab = None if labels is None else np.ascontiguousarray(labels, np.int64)
    out = np.zeros(len(pa), np.float32)
    _check(load().sphb_geodesic_hausdorff(n, None if rp is None else _p(rp, i64p), k, _p(ids, i64p), _p(dist, f32p), None if lab is None else _p(lab, i64p),
                                          len(comp_ptr) - 1, _p(comp_ptr, i64p), _p(comp_pix, i64p), _p(pa, i64p), _p(pb, i64p), len(pa), _p(out, f32p)))
    return out


class GeodesicGraph:
    """The data-level kNN graph resident on the device (sphb_geodesic_graph_create): upload once, ask for the distances of every
    level.  Use as a context manager or call close(); belongs to the thread that created it."""

    def __init__(self, ids, dist, rowptr=None, labels=None):
        n, rp, k, ids, dist = _geo_graph(ids, dist, rowptr)
        lab = None if labels is None else np.ascontiguousarray(labels, np.int64)
        self._h = C.c_void_p()
        _check(load().sphb_geodesic_graph_create(n, None if rp is None else _p(rp, i64p), k, _p(ids, i64p), _p(dist, f32p),
                                                 None if lab is None else _p(lab, i64p), C.byref(self._h)))
        self.n = n

    def distances(self, start, end) -> np.ndarray:
        start = np.ascontiguousarray(start, np.int64)
        end = np.ascontiguousarray(end, np.int64)
        out = np.zeros(len(start), np.float32)
        _check(load().sphb_geodesic_distances_g(self._h, _p(start, i64p), _p(end, i64p), len(start), _p(out, f32p)))
        return out

    def hausdorff(self, comp_ptr, comp_pix, pa, pb) -> np.ndarray:
        comp_ptr = np.ascontiguousarray(comp_ptr, np.int64)
        comp_pix = np.ascontiguousarray(comp_pix, np.int64)
        pa = np.ascontiguousarray(pa, np.int64)
        pb = np.ascontiguousarray(pb, np.int64)
        out = np.zeros(len(pa), np.float32)
        _check(load().sphb_geodesic_hausdorff_g(self._h, len(comp_ptr) - 1, _p(comp_ptr, i64p), _p(comp_pix, i64p), _p(pa, i64p), _p(pb, i64p), len(pa),
                                                _p(out, f32p)))
        return out

    def close(self):
        if self._h:
            load().sphb_geodesic_graph_destroy(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def geodesic_last_stats() -> dict:
    out = np.zeros(4, np.int64)
    _check(load().sphb_geodesic_last_stats(_p(out, i64p)))
    return dict(expanded=int(out[0]), passes=int(out[1]), searches=int(out[2]), sources=int(out[3]))


# ------------------------------------------------------------------ the reference's cache blobs (host code, no GPU needed)
def cache_chunk_limit(nbytes: int):
    """Byte limit above which the chunked layouts are written (INT_MAX in the reference; tests lower it).  0 restores the default."""
    load().sphb_debug_cache_chunk_limit(int(nbytes))


def lz4_compress(data: bytes) -> bytes:
    L = load()
    cap = int(L.sphb_lz4_compress_bound(len(data)))
    dst = C.create_string_buffer(cap)
    m = int(L.sphb_lz4_compress(data, len(data), dst, cap))
    if m < 0:
        raise SphbError("lz4_compress failed")
    return dst.raw[:m]


def lz4_decompress(block: bytes, size: int) -> bytes:
    dst = C.create_string_buffer(max(size, 1))
    m = int(load().sphb_lz4_decompress(block, len(block), dst, size))
    if m < 0:
        raise SphbError("lz4_decompress: malformed block or output too small")
    return dst.raw[:m]


def cache_write_vec(path, vec: np.ndarray):
    """writeCompressedVecToBinary<T> (sph/utils/FileIO.hpp:279-291)."""
    vec = np.ascontiguousarray(vec)
    _check(load().sphb_cache_write_vec(str(path).encode(), vec.ctypes.data_as(C.c_void_p), vec.size, vec.dtype.itemsize))


def cache_read_vec(path, dtype) -> np.ndarray:
    """loadCompressedVecFromBinary<T> (sph/utils/FileIO.hpp:523-640)."""
    dtype = np.dtype(dtype)
    cnt = np.zeros(1, np.int64)
    _check(load().sphb_cache_read_vec(str(path).encode(), None, 0, dtype.itemsize, _p(cnt, i64p)))
    out = np.zeros(max(int(cnt[0]), 1), dtype)
    _check(load().sphb_cache_read_vec(str(path).encode(), out.ctypes.data_as(C.c_void_p), len(out), dtype.itemsize, _p(cnt, i64p)))
    return out[: int(cnt[0])]


def cache_write_graph(base, ids, dist, rowptr=None, symmetric: bool = False):
    """writeCompressedGraphToBinary (sph/utils/FileIO.cpp:586-618): <base>{Distances,Indices,NNs,Symmetric}.cache."""
    n, rp, k, ids, dist = _geo_graph(ids, dist, rowptr)
    _check(load().sphb_cache_write_graph(str(base).encode(), n, None if rp is None else _p(rp, i64p), k, _p(ids, i64p), _p(dist, f32p), int(symmetric)))


def cache_read_graph(base):
    """loadCompressedGraphFromBinary (sph/utils/FileIO.cpp:895-937): (rowptr, ids, dist, symmetric)."""
    n = np.zeros(1, np.int64)
    nnz = np.zeros(1, np.int64)
    _check(load().sphb_cache_read_graph(str(base).encode(), 0, 0, _p(n, i64p), _p(nnz, i64p), None, None, None, None))
    nns = np.zeros(max(int(n[0]), 1), np.int64)
    ids = np.zeros(max(int(nnz[0]), 1), np.int64)
    dist = np.zeros(max(int(nnz[0]), 1), np.float32)
    sym = C.c_int(0)
    _check(load().sphb_cache_read_graph(str(base).encode(), len(nns), len(ids), _p(n, i64p), _p(nnz, i64p), _p(nns, i64p), _p(ids, i64p), _p(dist, f32p),
                                        C.byref(sym)))
    rowptr = np.concatenate([[0], np.cumsum(nns[: int(n[0])])]).astype(np.int64)
    return rowptr, ids[: int(nnz[0])], dist[: int(nnz[0])], bool(sym.value)


def cache_write_prob(path, rowptr, col, val):
    """writeCompressedSparseMatHDIToBinary (sph/utils/FileIO.cpp:620-799)."""
    rowptr = np.ascontiguousarray(rowptr, np.int64)
    col = np.ascontiguousarray(col, np.uint32)
    val = np.ascontiguousarray(val, np.float32)
    _check(load().sphb_cache_write_prob(str(path).encode(), len(rowptr) - 1, _p(rowptr, i64p), _p(col, u32p), _p(val, f32p)))


def cache_read_prob(path):
    """loadCompressedSparseMatHDIFromBinary (sph/utils/FileIO.cpp:1274-1460): (rowptr, col, val)."""
    n = np.zeros(1, np.int64)
    nnz = np.zeros(1, np.int64)
    _check(load().sphb_cache_read_prob(str(path).encode(), 0, 0, _p(n, i64p), _p(nnz, i64p), None, None, None))
    rowptr = np.zeros(int(n[0]) + 1, np.int64)
    col = np.zeros(max(int(nnz[0]), 1), np.uint32)
    val = np.zeros(max(int(nnz[0]), 1), np.float32)
    _check(load().sphb_cache_read_prob(str(path).encode(), int(n[0]), len(col), _p(n, i64p), _p(nnz, i64p), _p(rowptr, i64p), _p(col, u32p), _p(val, f32p)))
    return rowptr, col[: int(nnz[0])], val[: int(nnz[0])]


# ------------------------------------------------------------------ resident level pipeline
def level_prob_dist(rowptr, col, val, k: int, prune_val: float = 1e-4, variant: int = 0, comp_sizes=None, pair_sims: bool = True,
                    perplexity: float = 0.0, scheme: int = 1, symmetrize: bool = True):
    """LevelSimilarities' useRandomWalks branch + symmetrizeOutput in one call (host buffers): returns the level's P as CSR."""
    rowptr = np.ascontiguousarray(rowptr, np.int64)
    col = np.ascontiguousarray(col, np.uint32)
    val = np.ascontiguousarray(val, np.float32)
    n = len(rowptr) - 1
    cap = max(1, 2 * n * k)
    orp = np.zeros(n + 1, np.int64)
    oc = np.zeros(cap, np.uint32)
    ov = np.zeros(cap, np.float32)
    nnz = np.zeros(1, np.int64)
    cs = None if comp_sizes is None else np.ascontiguousarray(comp_sizes, np.int64)
    _check(load().sphb_level_prob_dist(n, _p(rowptr, i64p), _p(col, u32p), _p(val, f32p), k, prune_val, variant, None if cs is None else _p(cs, i64p),
                                       int(pair_sims), perplexity, scheme, int(symmetrize), _p(orp, i64p), _p(oc, u32p), _p(ov, f32p), cap,
                                       _p(nnz, i64p)))
    m = int(nnz[0])
    return orp, oc[:m].copy(), ov[:m].copy()


def level_last_timing() -> dict:
    out = np.zeros(5, np.float32)
    _check(load().sphb_level_last_timing(_p(out, f32p)))
    return dict(similarities_ms=float(out[0]), to_csr_ms=float(out[1]), gaussian_ms=float(out[2]), remove_zeros_ms=float(out[3]),
                symmetrize_ms=float(out[4]))


def random_walks_dev(n, rowptr_ptr, col_ptr, val_ptr, num_walks, walk_len, weighting, seed, q0, q1, out_rowptr_ptr, out_col_ptr, out_val_ptr, cap,
                     stream, prune_value=0.0, prune_steps=0, remove_diagonal=True, normalize=True) -> int:
    nnz = np.zeros(1, np.int64)
    _check(load().sphb_random_walks_dev(n, rowptr_ptr, col_ptr, val_ptr, num_walks, walk_len, weighting, prune_value, prune_steps, seed,
                                        int(remove_diagonal), int(normalize), q0, q1, out_rowptr_ptr, out_col_ptr, out_val_ptr, cap, _p(nnz, i64p), stream))
    return int(nnz[0])


def rw_merge_dev(n, rowptr_ptr, col_ptr, val_ptr, nnz_in, parents_ptr, n_merged, norm, weight_by_size, out_rowptr_ptr, out_col_ptr, out_val_ptr, cap,
                 stream) -> int:
    nnz = np.zeros(1, np.int64)
    _check(load().sphb_rw_merge_dev(n, rowptr_ptr, col_ptr, val_ptr, nnz_in, parents_ptr, n_merged, int(norm), int(weight_by_size), out_rowptr_ptr,
                                    out_col_ptr, out_val_ptr, cap, _p(nnz, i64p), stream))
    return int(nnz[0])


def create_similarities_dev(n, rowptr_ptr, col_ptr, val_ptr, nnz_in, k, prune_val, variant, comp_sizes_ptr, out_cols_ptr, out_vals_ptr, row_nnz_ptr, stream):
    _check(load().sphb_create_similarities_dev(n, rowptr_ptr, col_ptr, val_ptr, nnz_in, k, prune_val, variant, comp_sizes_ptr, out_cols_ptr, out_vals_ptr,
                                               row_nnz_ptr, stream))


def create_similarities_slab_dev(n, rowptr_ptr, col_ptr, val_ptr, nnz_in, k, prune_val, variant, comp_sizes_ptr, r0, r1, out_cols_ptr, out_vals_ptr,
                                 row_nnz_ptr, stream):
    _check(load().sphb_create_similarities_slab_dev(n, rowptr_ptr, col_ptr, val_ptr, nnz_in, k, prune_val, variant, comp_sizes_ptr, r0, r1, out_cols_ptr,
                                                    out_vals_ptr, row_nnz_ptr, stream))


def level_from_sims_dev(n, k, cols_ptr, vals_ptr, row_nnz_ptr, perplexity, scheme, symmetrize, out_rowptr_ptr, out_col_ptr, out_val_ptr, cap, stream) -> int:
    nnz = np.zeros(1, np.int64)
    _check(load().sphb_level_from_sims_dev(n, k, cols_ptr, vals_ptr, row_nnz_ptr, perplexity, scheme, int(symmetrize), out_rowptr_ptr, out_col_ptr,
                                           out_val_ptr, cap, _p(nnz, i64p), stream))
    return int(nnz[0])


def prow_gaussian_inplace_dev(n, rowptr_ptr, val_ptr, max_row_nnz, perplexity, ignore, stream):
    _check(load().sphb_prow_gaussian_inplace_dev(n, rowptr_ptr, val_ptr, max_row_nnz, perplexity, ignore, stream))


def symmetrize_dev(n, rowptr_ptr, col_ptr, val_ptr, nnz_in, umap, out_rowptr_ptr, out_col_ptr, out_val_ptr, cap, stream) -> int:
    nnz = np.zeros(1, np.int64)
    _check(load().sphb_symmetrize_dev(n, rowptr_ptr, col_ptr, val_ptr, nnz_in, int(umap), out_rowptr_ptr, out_col_ptr, out_val_ptr, cap, _p(nnz, i64p),
                                      stream))
    return int(nnz[0])


def find_k_dev(n, rowptr_ptr, col_ptr, val_ptr, max_row_nnz, k, top, out_cols_ptr, out_vals_ptr, row_nnz_ptr, stream):
    _check(load().sphb_find_k_dev(n, rowptr_ptr, col_ptr, val_ptr, max_row_nnz, k, int(top), out_cols_ptr, out_vals_ptr, row_nnz_ptr, stream))


def level_prob_dist_dev(n, rowptr_ptr, col_ptr, val_ptr, nnz_in, k, prune_val, variant, comp_sizes_ptr, pair_sims, perplexity, scheme, symmetrize,
                        out_rowptr_ptr, out_col_ptr, out_val_ptr, cap, stream) -> int:
    nnz = np.zeros(1, np.int64)
    _check(load().sphb_level_prob_dist_dev(n, rowptr_ptr, col_ptr, val_ptr, nnz_in, k, prune_val, variant, comp_sizes_ptr, int(pair_sims), perplexity,
                                           scheme, int(symmetrize), out_rowptr_ptr, out_col_ptr, out_val_ptr, cap, _p(nnz, i64p), stream))
    return int(nnz[0])


def knn_rows_noalloc_dev(x_ptr: int, n: int, d: int, k: int, metric: int, l2_squared: bool, rows_ptr: int, nrows: int, ids_ptr: int, dist_ptr: int,
                         stream: int):
    """Exact kNN of a list of <= 256 query rows (k <= 128) with the library's fixed scratch: no device allocation."""
    _check(load().sphb_knn_rows_noalloc_dev(x_ptr, n, d, k, metric, int(l2_squared), rows_ptr, nrows, ids_ptr, dist_ptr, stream))


def renorm_l2(x: np.ndarray) -> np.ndarray:
    """COSINE preparation (faiss::fvec_renorm_L2 on a copy, sph/NearestNeighbors.cpp:208-222) on the device."""
    x = np.ascontiguousarray(x, np.float32)
    out = np.empty_like(x)
    _check(load().sphb_renorm_l2(_p(x, f32p), x.shape[0], x.shape[1], _p(out, f32p)))
    return out


def rw_post_merge(rowptr, col, val, remove_diagonal: bool, norm: int):
    """Hierarchy::updateRandomWalks after the merge: diagonal / zero removal, then norm 0 none, 1 ONEDIM, 2 TWODIM."""
    rowptr = np.ascontiguousarray(rowptr, np.int64)
    col = np.ascontiguousarray(col, np.uint32)
    val = np.ascontiguousarray(val, np.float32)
    n = len(rowptr) - 1
    cap = max(1, len(col))
    orp = np.zeros(n + 1, np.int64)
    oc = np.zeros(cap, np.uint32)
    ov = np.zeros(cap, np.float32)
    nnz = np.zeros(1, np.int64)
    _check(load().sphb_rw_post_merge(n, _p(rowptr, i64p), _p(col, u32p), _p(val, f32p), int(remove_diagonal), int(norm), _p(orp, i64p), _p(oc, u32p),
                                     _p(ov, f32p), _p(nnz, i64p)))
    m = int(nnz[0])
    return orp, oc[:m].copy(), ov[:m].copy()
