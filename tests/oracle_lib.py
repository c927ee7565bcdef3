"""ctypes view of the CPU oracle (oracle/_build/libsph_oracle.so) — TEST INFRASTRUCTURE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs import this module.
"""
from __future__ import annotations

import ctypes as C
import glob
import os
import subprocess
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
ORACLE_DIR = ROOT / "oracle"
LIB = ORACLE_DIR / "_build" / "libsph_oracle.so"

i64p = C.POINTER(C.c_int64)
u64p = C.POINTER(C.c_uint64)
u32p = C.POINTER(C.c_uint32)
i32p = C.POINTER(C.c_int32)
f32p = C.POINTER(C.c_float)
f64p = C.POINTER(C.c_double)


def _p(a, t):
    return None if a is None else a.ctypes.data_as(t)


def build():
    subprocess.run(["make", "-C", str(ORACLE_DIR)], check=True, stdout=subprocess.DEVNULL)


class Oracle:
    def __init__(self, lib):
        self.lib = lib
        L = lib
        L.orc_knn_exact.argtypes = [f32p, C.c_int64, C.c_int, C.c_int, C.c_int, i64p, f32p]
        L.orc_knn_exact_rows.argtypes = [f32p, C.c_int64, C.c_int, C.c_int, C.c_int, i64p, C.c_int64, i64p, f32p]
        L.orc_renorm_l2.argtypes = [f32p, C.c_int64, C.c_int, f32p]
        L.orc_knn_postprocess.argtypes = [i64p, f32p, C.c_int64, C.c_int, C.c_int]
        L.orc_knn_postprocess.restype = C.c_int64
        L.orc_knn_postprocess_rows.argtypes = [i64p, f32p, C.c_int64, C.c_int, C.c_int, i64p]
        L.orc_knn_postprocess_rows.restype = C.c_int64
        L.orc_check_all_neighbors_exist.argtypes = [i64p, C.c_int64]
        L.orc_blas_init.argtypes = [C.c_char_p]
        L.orc_knn_blas_l2.argtypes = [f32p, C.c_int64, C.c_int, C.c_int, C.c_int64, C.c_int64, i64p, f32p]
        L.orc_knn_blas_last_stats.argtypes = [f64p]
        L.orc_knn_blas_last_stats.restype = None
        L.orc_sgemm_standalone_gflops.argtypes = [C.c_int, C.c_double]
        L.orc_sgemm_standalone_gflops.restype = C.c_double
        L.orc_gaussian_row.argtypes = [f32p, f32p, C.c_int, C.c_double, C.c_int, C.c_double, C.c_int]
        L.orc_gaussian_row.restype = C.c_double
        L.orc_prow_gaussian_graph.argtypes = [i64p, f32p, C.c_int64, C.c_int, C.c_float, u32p, f32p, f64p]
        L.orc_prow_gaussian_inplace.argtypes = [C.c_int64, i64p, f32p, C.c_float, C.c_int]
        L.orc_prow_gaussian_flat.argtypes = [f32p, C.c_int64, C.c_int, C.c_double, f32p]
        L.orc_prow_linear_graph.argtypes = [i64p, f32p, C.c_int64, C.c_int, u32p, f32p]
        L.orc_prow_exponential_graph.argtypes = [i64p, f32p, C.c_int64, C.c_int, C.c_int, u32p, f32p]
        L.orc_prow_exponential_inplace.argtypes = [C.c_int64, i64p, u32p, f32p, C.c_int]
        L.orc_symmetrize.argtypes = [C.c_int64, i64p, u32p, f32p, C.c_int, i64p, u32p, f32p, C.c_int64]
        L.orc_symmetrize.restype = C.c_int64
        L.orc_level_schedule.argtypes = [i64p, C.c_int, C.c_int64, f32p, i64p]
        L.orc_random_walks.argtypes = [C.c_int64, i64p, u32p, f32p, C.c_uint64, C.c_uint64, C.c_int, C.c_float, C.c_uint64, C.c_uint64, C.c_int,
                                       C.c_int, C.c_int, i64p, u32p, f32p, C.c_int64]
        L.orc_random_walks.restype = C.c_int64
        L.orc_rw_merge.argtypes = [C.c_int64, i64p, u32p, f32p, u64p, C.c_int64, C.c_int, C.c_int, i64p, u32p, f32p, C.c_int64]
        L.orc_rw_merge.restype = C.c_int64
        L.orc_create_similarities.argtypes = [C.c_int64, i64p, u32p, f32p, C.c_int64, C.c_float, C.c_int, i64p, u32p, f32p, i32p]
        L.orc_create_similarities_mergedot.argtypes = [C.c_int64, i64p, u32p, f32p, C.c_int64, C.c_float, C.c_int, i64p, u32p, f32p, i32p]
        L.orc_create_similarities_rows.argtypes = [C.c_int64, i64p, u32p, f32p, C.c_int64, C.c_float, C.c_int, i64p, C.c_int64, C.c_int64, u32p, f32p, i32p]
        L.orc_create_similarities_sph.argtypes = [C.c_int64, i64p, u32p, f32p, C.c_float, i64p, u32p, f32p, C.c_int64]
        L.orc_create_similarities_sph.restype = C.c_int64
        L.orc_find_k.argtypes = [u32p, f32p, C.c_int, C.c_int, C.c_int, u32p, f32p]
        L.orc_pair_sims.argtypes = [C.c_int64, i64p, u32p, f32p, i64p, i64p, C.c_int64, C.c_int, f32p]
        L.orc_neighbor_overlap.argtypes = [i64p, C.c_int, i64p, i64p, i64p, i64p, C.c_int64, f32p]
        L.orc_symmetrize_graph.argtypes = [C.c_int64, i64p, i64p, f32p, i64p, i64p, f32p, C.c_int64]
        L.orc_symmetrize_graph.restype = C.c_int64
        L.orc_merge_graph_nodes.argtypes = [C.c_int64, i64p, i64p, f32p, u64p, C.c_int64, i64p, i64p, f32p, C.c_int64]
        L.orc_merge_graph_nodes.restype = C.c_int64
        L.orc_merge_data_distances.argtypes = [C.c_int64, i64p, u32p, f32p, u64p, C.c_int64, i64p, u32p, f32p, C.c_int64]
        L.orc_merge_data_distances.restype = C.c_int64
        L.orc_rw_as_knn_distances.argtypes = [C.c_int64, i64p, u32p, f32p, i64p, i64p, f32p, C.c_int64]
        L.orc_rw_as_knn_distances.restype = C.c_int64
        L.orc_remove_zeros.argtypes = [C.c_int64, i64p, u32p, f32p, i64p, u32p, f32p]
        L.orc_remove_zeros.restype = C.c_int64
        L.orc_euclid_hausdorff.argtypes = [f32p, C.c_int, i64p, i64p, i64p, i64p, C.c_int64, f32p]
        L.orc_prow_graph_ragged.argtypes = [C.c_int64, i64p, i64p, f32p, C.c_float, C.c_int, i64p, u32p, f32p]
        L.orc_weak_components.argtypes = [C.c_int64, i64p, i64p, i64p]
        L.orc_weak_components.restype = C.c_int64
        L.orc_component_centers.argtypes = [f32p, C.c_int64, C.c_int, i64p, C.c_int64, f32p]
        L.orc_bridge_components.argtypes = [f32p, C.c_int64, C.c_int, i64p, i64p, C.c_int64, i64p, i64p, f32p]
        L.orc_level_knn_overlap.argtypes = [i64p, C.c_int, C.c_int64, i64p, i64p, C.c_int64, i64p, f32p]
        L.orc_level_knn_walks.argtypes = [C.c_int64, i64p, u32p, f32p, C.c_int, C.c_int64, i64p, f32p]
        L.orc_shortest_paths.argtypes = [C.c_int64, i64p, i64p, f32p, i64p, C.c_int64, i64p, i64p, f32p]
        L.orc_sssp.argtypes = [C.c_int64, i64p, i64p, f32p, C.c_int64, f32p]
        L.orc_astar.argtypes = [C.c_int64, C.c_int, i64p, f32p, f32p, C.c_int, C.c_int64, i64p, i64p, f32p]
        L.orc_geodesic_hausdorff.argtypes = [C.c_int64, i64p, i64p, f32p, i64p, i64p, i64p, i64p, i64p, C.c_int64, f32p]

    # ---- Stage A
    def knn_exact(self, x, k, metric=1, rows=None):
        """metric: faiss values, 1 = L2 (squared distances), 0 = inner product."""
        x = np.ascontiguousarray(x, np.float32)
        n, d = x.shape
        if rows is None:
            ids = np.empty((n, k), np.int64)
            dist = np.empty((n, k), np.float32)
            self.lib.orc_knn_exact(_p(x, f32p), n, d, k, metric, _p(ids, i64p), _p(dist, f32p))
        else:
            rows = np.ascontiguousarray(rows, np.int64)
            ids = np.empty((len(rows), k), np.int64)
            dist = np.empty((len(rows), k), np.float32)
            self.lib.orc_knn_exact_rows(_p(x, f32p), n, d, k, metric, _p(rows, i64p), len(rows), _p(ids, i64p), _p(dist, f32p))
        return ids, dist

    def knn_postprocess(self, ids, dist, l2_squared=False, row_ids=None):
        """In place; rows are assumed to be points 0..n-1 unless row_ids gives the point id of each row."""
        n, k = ids.shape
        if row_ids is None:
            return int(self.lib.orc_knn_postprocess(_p(ids, i64p), _p(dist, f32p), n, k, int(l2_squared)))
        row_ids = np.ascontiguousarray(row_ids, np.int64)
        return int(self.lib.orc_knn_postprocess_rows(_p(ids, i64p), _p(dist, f32p), n, k, int(l2_squared), _p(row_ids, i64p)))

    def knn_reference_graph(self, x, k, l2_squared=False, rows=None, metric=1):
        ids, dist = self.knn_exact(x, k, metric, rows)
        self.knn_postprocess(ids, dist, l2_squared, rows)
        return ids, dist

    def blas_init(self):
        import numpy

        cands = glob.glob(os.path.join(os.path.dirname(numpy.__file__), "..", "numpy.libs", "libscipy_openblas64_*.so"))
        if not cands:
            return False
        return self.lib.orc_blas_init(os.path.abspath(cands[0]).encode()) == 0

    def knn_blas_l2(self, x, k, q0, q1):
        x = np.ascontiguousarray(x, np.float32)
        n, d = x.shape
        ids = np.empty((q1 - q0, k), np.int64)
        dist = np.empty((q1 - q0, k), np.float32)
        rc = self.lib.orc_knn_blas_l2(_p(x, f32p), n, d, k, q0, q1, _p(ids, i64p), _p(dist, f32p))
        if rc != 0:
            raise RuntimeError("orc_knn_blas_l2 failed (OpenBLAS not initialised)")
        return ids, dist

    def knn_blas_last_stats(self):
        out = np.zeros(4, np.float64)
        self.lib.orc_knn_blas_last_stats(_p(out, f64p))
        return dict(sgemm_s=float(out[0]), heap_s=float(out[1]), sgemm_flop=float(out[2]), threads=int(out[3]),
                    sgemm_gflops=float(out[2] / out[0] / 1e9) if out[0] > 0 else 0.0)

    def sgemm_standalone_gflops(self, d, seconds=1.0):
        return float(self.lib.orc_sgemm_standalone_gflops(int(d), float(seconds)))

    # ---- Stage B
    def gaussian_row(self, dist, perplexity, tol=1e-6, ignore=0, max_iter=200):
        dist = np.ascontiguousarray(dist, np.float32)
        prob = np.zeros_like(dist)
        sigma = self.lib.orc_gaussian_row(_p(dist, f32p), _p(prob, f32p), len(dist), perplexity, max_iter, tol, ignore)
        return prob, sigma

    def prow_gaussian_graph(self, ids, dist, perplexity=0.0):
        ids = np.ascontiguousarray(ids, np.int64)
        dist = np.ascontiguousarray(dist, np.float32)
        n, k = ids.shape
        cols = np.zeros((n, k - 1), np.uint32)
        vals = np.zeros((n, k - 1), np.float32)
        sigma = np.zeros(n, np.float64)
        self.lib.orc_prow_gaussian_graph(_p(ids, i64p), _p(dist, f32p), n, k, perplexity, _p(cols, u32p), _p(vals, f32p), _p(sigma, f64p))
        return cols, vals, sigma

    def prow_gaussian_inplace(self, rowptr, val, perplexity=0.0, ignore=-1):
        rowptr = np.ascontiguousarray(rowptr, np.int64)
        out = np.array(val, np.float32, copy=True)
        self.lib.orc_prow_gaussian_inplace(len(rowptr) - 1, _p(rowptr, i64p), _p(out, f32p), perplexity, ignore)
        return out

    def prow_gaussian_flat(self, dist, perplexity):
        dist = np.ascontiguousarray(dist, np.float32)
        n, k = dist.shape
        temp = np.zeros((n, k), np.float32)
        self.lib.orc_prow_gaussian_flat(_p(dist, f32p), n, k, perplexity, _p(temp, f32p))
        return temp

    def prow_exponential_graph(self, ids, dist, sort_cols=False):
        ids = np.ascontiguousarray(ids, np.int64)
        dist = np.ascontiguousarray(dist, np.float32)
        n, k = ids.shape
        cols = np.zeros((n, k - 1), np.uint32)
        vals = np.zeros((n, k - 1), np.float32)
        self.lib.orc_prow_exponential_graph(_p(ids, i64p), _p(dist, f32p), n, k, int(sort_cols), _p(cols, u32p), _p(vals, f32p))
        return cols, vals

    def prow_exponential_inplace(self, rowptr, val, col=None, ignore=-1):
        rowptr = np.ascontiguousarray(rowptr, np.int64)
        out = np.array(val, np.float32, copy=True)
        colp = _p(np.ascontiguousarray(col, np.uint32), u32p) if col is not None else None
        self.lib.orc_prow_exponential_inplace(len(rowptr) - 1, _p(rowptr, i64p), colp, _p(out, f32p), ignore)
        return out

    def prow_linear_graph(self, ids, dist):
        ids = np.ascontiguousarray(ids, np.int64)
        dist = np.ascontiguousarray(dist, np.float32)
        n, k = ids.shape
        cols = np.zeros((n, k - 1), np.uint32)
        vals = np.zeros((n, k - 1), np.float32)
        self.lib.orc_prow_linear_graph(_p(ids, i64p), _p(dist, f32p), n, k, _p(cols, u32p), _p(vals, f32p))
        return cols, vals

    def symmetrize(self, rowptr, col, val, umap=False):
        rowptr = np.ascontiguousarray(rowptr, np.int64)
        col = np.ascontiguousarray(col, np.uint32)
        val = np.ascontiguousarray(val, np.float32)
        n = len(rowptr) - 1
        cap = max(1, 2 * len(col))
        orp = np.zeros(n + 1, np.int64)
        oc = np.zeros(cap, np.uint32)
        ov = np.zeros(cap, np.float32)
        m = self.lib.orc_symmetrize(n, _p(rowptr, i64p), _p(col, u32p), _p(val, f32p), int(umap), _p(orp, i64p), _p(oc, u32p), _p(ov, f32p), cap)
        assert m >= 0
        return orp, oc[:m].copy(), ov[:m].copy()

    def level_schedule(self, level_sizes, k0):
        ls = np.ascontiguousarray(level_sizes, np.int64)
        perp = np.zeros(len(ls), np.float32)
        ks = np.zeros(len(ls), np.int64)
        self.lib.orc_level_schedule(_p(ls, i64p), len(ls), k0, _p(perp, f32p), _p(ks, i64p))
        return perp, ks

    # ---- Stage C
    def random_walks(self, rowptr, col, val, num_walks, walk_len, weighting=0, prune_value=0.0, prune_steps=0, seed=0, threads=1,
                     remove_diagonal=True, normalize=True):
        rowptr = np.ascontiguousarray(rowptr, np.int64)
        col = np.ascontiguousarray(col, np.uint32)
        val = np.ascontiguousarray(val, np.float32)
        n = len(rowptr) - 1
        cap = max(1, n * min(n, num_walks * walk_len))
        orp = np.zeros(n + 1, np.int64)
        oc = np.zeros(cap, np.uint32)
        ov = np.zeros(cap, np.float32)
        m = self.lib.orc_random_walks(n, _p(rowptr, i64p), _p(col, u32p), _p(val, f32p), num_walks, walk_len, weighting, prune_value, prune_steps,
                                      seed, threads, int(remove_diagonal), int(normalize), _p(orp, i64p), _p(oc, u32p), _p(ov, f32p), cap)
        assert m >= 0
        return orp, oc[:m].copy(), ov[:m].copy()

    def rw_merge(self, rowptr, col, val, parents, n_merged, norm=True, weight_by_size=True):
        rowptr = np.ascontiguousarray(rowptr, np.int64)
        col = np.ascontiguousarray(col, np.uint32)
        val = np.ascontiguousarray(val, np.float32)
        parents = np.ascontiguousarray(parents, np.uint64)
        n = len(rowptr) - 1
        cap = max(1, len(col))
        orp = np.zeros(n_merged + 1, np.int64)
        oc = np.zeros(cap, np.uint32)
        ov = np.zeros(cap, np.float32)
        m = self.lib.orc_rw_merge(n, _p(rowptr, i64p), _p(col, u32p), _p(val, f32p), _p(parents, u64p), n_merged, int(norm), int(weight_by_size),
                                  _p(orp, i64p), _p(oc, u32p), _p(ov, f32p), cap)
        assert m >= 0
        return orp, oc[:m].copy(), ov[:m].copy()

    # ---- Stage D
    def create_similarities(self, rowptr, col, val, k, prune_val=1e-4, variant=0, comp_sizes=None):
        rowptr = np.ascontiguousarray(rowptr, np.int64)
        col = np.ascontiguousarray(col, np.uint32)
        val = np.ascontiguousarray(val, np.float32)
        n = len(rowptr) - 1
        oc = np.zeros((n, k), np.uint32)
        ov = np.zeros((n, k), np.float32)
        nnz = np.zeros(n, np.int32)
        cs = None if comp_sizes is None else np.ascontiguousarray(comp_sizes, np.int64)
        self.lib.orc_create_similarities(n, _p(rowptr, i64p), _p(col, u32p), _p(val, f32p), k, prune_val, variant, _p(cs, i64p), _p(oc, u32p),
                                         _p(ov, f32p), _p(nnz, i32p))
        return oc, ov, nnz

    def create_similarities_rows(self, rowptr, col, val, k, row0, row1, prune_val=1e-4, variant=0, comp_sizes=None, mergedot=False):
        """rows [row0, row1) only (the bounded CPU-baseline sample); mergedot=True: the first restatement, all rows."""
        rowptr = np.ascontiguousarray(rowptr, np.int64)
        col = np.ascontiguousarray(col, np.uint32)
        val = np.ascontiguousarray(val, np.float32)
        n = len(rowptr) - 1
        if mergedot:
            row0, row1 = 0, n
        m = row1 - row0
        oc = np.zeros((m, k), np.uint32)
        ov = np.zeros((m, k), np.float32)
        nnz = np.zeros(m, np.int32)
        cs = None if comp_sizes is None else np.ascontiguousarray(comp_sizes, np.int64)
        csp = _p(cs, i64p) if cs is not None else None
        if mergedot:
            self.lib.orc_create_similarities_mergedot(n, _p(rowptr, i64p), _p(col, u32p), _p(val, f32p), k, prune_val, variant, csp, _p(oc, u32p),
                                                      _p(ov, f32p), _p(nnz, i32p))
        else:
            self.lib.orc_create_similarities_rows(n, _p(rowptr, i64p), _p(col, u32p), _p(val, f32p), k, prune_val, variant, csp, row0, row1,
                                                  _p(oc, u32p), _p(ov, f32p), _p(nnz, i32p))
        return oc, ov, nnz

    def create_similarities_sph(self, rowptr, col, val, prune_val=1e-4):
        rowptr = np.ascontiguousarray(rowptr, np.int64)
        col = np.ascontiguousarray(col, np.uint32)
        val = np.ascontiguousarray(val, np.float32)
        n = len(rowptr) - 1
        cap = n * n
        orp = np.zeros(n + 1, np.int64)
        oc = np.zeros(cap, np.uint32)
        ov = np.zeros(cap, np.float32)
        m = self.lib.orc_create_similarities_sph(n, _p(rowptr, i64p), _p(col, u32p), _p(val, f32p), prune_val, _p(orp, i64p), _p(oc, u32p), _p(ov, f32p), cap)
        assert m >= 0
        return orp, oc[:m].copy(), ov[:m].copy()

    def find_k(self, col, val, k, top):
        col = np.ascontiguousarray(col, np.uint32)
        val = np.ascontiguousarray(val, np.float32)
        oc = np.zeros(k, np.uint32)
        ov = np.zeros(k, np.float32)
        m = self.lib.orc_find_k(_p(col, u32p), _p(val, f32p), len(col), k, int(top), _p(oc, u32p), _p(ov, f32p))
        return oc[:m].copy(), ov[:m].copy()

    def pair_sims(self, rowptr, col, val, pa, pb, mode=0):
        rowptr = np.ascontiguousarray(rowptr, np.int64)
        col = np.ascontiguousarray(col, np.uint32)
        val = np.ascontiguousarray(val, np.float32)
        pa = np.ascontiguousarray(pa, np.int64)
        pb = np.ascontiguousarray(pb, np.int64)
        out = np.zeros(len(pa), np.float32)
        self.lib.orc_pair_sims(len(rowptr) - 1, _p(rowptr, i64p), _p(col, u32p), _p(val, f32p), _p(pa, i64p), _p(pb, i64p), len(pa), mode, _p(out, f32p))
        return out

    def neighbor_overlap(self, knn_ids, comp_ptr, comp_pix, pa, pb):
        knn_ids = np.ascontiguousarray(knn_ids, np.int64)
        comp_ptr = np.ascontiguousarray(comp_ptr, np.int64)
        comp_pix = np.ascontiguousarray(comp_pix, np.int64)
        pa = np.ascontiguousarray(pa, np.int64)
        pb = np.ascontiguousarray(pb, np.int64)
        out = np.zeros(len(pa), np.float32)
        self.lib.orc_neighbor_overlap(_p(knn_ids, i64p), knn_ids.shape[1], _p(comp_ptr, i64p), _p(comp_pix, i64p), _p(pa, i64p), _p(pb, i64p), len(pa), _p(out, f32p))
        return out


    # ---- ragged-graph passes
    @staticmethod
    def _graph(ids, dist, rowptr):
        ids = np.ascontiguousarray(ids, np.int64)
        dist = np.ascontiguousarray(dist, np.float32)
        if rowptr is None:
            n, k = ids.shape
            rowptr = np.arange(n + 1, dtype=np.int64) * k
        return np.ascontiguousarray(rowptr, np.int64), ids.reshape(-1), dist.reshape(-1)

    def symmetrize_graph(self, ids, dist, rowptr=None):
        rowptr, ids, dist = self._graph(ids, dist, rowptr)
        n = len(rowptr) - 1
        cap = max(1, 2 * len(ids))
        orp = np.zeros(n + 1, np.int64)
        oi = np.zeros(cap, np.int64)
        od = np.zeros(cap, np.float32)
        m = self.lib.orc_symmetrize_graph(n, _p(rowptr, i64p), _p(ids, i64p), _p(dist, f32p), _p(orp, i64p), _p(oi, i64p), _p(od, f32p), cap)
        assert m >= 0
        return orp, oi[:m].copy(), od[:m].copy()

    def merge_graph_nodes(self, ids, dist, parents, n_merged, rowptr=None):
        rowptr, ids, dist = self._graph(ids, dist, rowptr)
        parents = np.ascontiguousarray(parents, np.uint64)
        n = len(rowptr) - 1
        cap = max(1, len(ids))
        orp = np.zeros(n_merged + 1, np.int64)
        oi = np.zeros(cap, np.int64)
        od = np.zeros(cap, np.float32)
        m = self.lib.orc_merge_graph_nodes(n, _p(rowptr, i64p), _p(ids, i64p), _p(dist, f32p), _p(parents, u64p), n_merged, _p(orp, i64p),
                                           _p(oi, i64p), _p(od, f32p), cap)
        assert m >= 0
        return orp, oi[:m].copy(), od[:m].copy()

    def merge_data_distances(self, rowptr, col, val, parents, n_merged):
        rowptr = np.ascontiguousarray(rowptr, np.int64)
        col = np.ascontiguousarray(col, np.uint32)
        val = np.ascontiguousarray(val, np.float32)
        parents = np.ascontiguousarray(parents, np.uint64)
        n = len(rowptr) - 1
        cap = max(1, len(col))
        orp = np.zeros(n_merged + 1, np.int64)
        oc = np.zeros(cap, np.uint32)
        ov = np.zeros(cap, np.float32)
        m = self.lib.orc_merge_data_distances(n, _p(rowptr, i64p), _p(col, u32p), _p(val, f32p), _p(parents, u64p), n_merged, _p(orp, i64p),
                                              _p(oc, u32p), _p(ov, f32p), cap)
        assert m >= 0
        return orp, oc[:m].copy(), ov[:m].copy()

    def rw_as_knn_distances(self, rowptr, col, val):
        rowptr = np.ascontiguousarray(rowptr, np.int64)
        col = np.ascontiguousarray(col, np.uint32)
        val = np.ascontiguousarray(val, np.float32)
        n = len(rowptr) - 1
        cap = max(1, len(col))
        orp = np.zeros(n + 1, np.int64)
        oi = np.zeros(cap, np.int64)
        od = np.zeros(cap, np.float32)
        m = self.lib.orc_rw_as_knn_distances(n, _p(rowptr, i64p), _p(col, u32p), _p(val, f32p), _p(orp, i64p), _p(oi, i64p), _p(od, f32p), cap)
        assert m >= 0
        return orp, oi[:m].copy(), od[:m].copy()

    def remove_zeros(self, rowptr, col, val):
        rowptr = np.ascontiguousarray(rowptr, np.int64)
        col = np.ascontiguousarray(col, np.uint32)
        val = np.ascontiguousarray(val, np.float32)
        n = len(rowptr) - 1
        cap = max(1, len(col))
        orp = np.zeros(n + 1, np.int64)
        oc = np.zeros(cap, np.uint32)
        ov = np.zeros(cap, np.float32)
        m = self.lib.orc_remove_zeros(n, _p(rowptr, i64p), _p(col, u32p), _p(val, f32p), _p(orp, i64p), _p(oc, u32p), _p(ov, f32p))
        return orp, oc[:m].copy(), ov[:m].copy()

    def euclid_hausdorff(self, x, comp_ptr, comp_pix, pa, pb):
        x = np.ascontiguousarray(x, np.float32)
        comp_ptr = np.ascontiguousarray(comp_ptr, np.int64)
        comp_pix = np.ascontiguousarray(comp_pix, np.int64)
        pa = np.ascontiguousarray(pa, np.int64)
        pb = np.ascontiguousarray(pb, np.int64)
        out = np.zeros(len(pa), np.float32)
        self.lib.orc_euclid_hausdorff(_p(x, f32p), x.shape[1], _p(comp_ptr, i64p), _p(comp_pix, i64p), _p(pa, i64p), _p(pb, i64p), len(pa), _p(out, f32p))
        return out

    # ---- 8(f)-4 geodesic distances (oracle_geodesic.cpp); graphs as (rowptr, ids, dist), rowptr None = fixed k from ids.shape
    @staticmethod
    def _graph_csr(rowptr, ids, dist):
        ids = np.ascontiguousarray(ids, np.int64)
        dist = np.ascontiguousarray(dist, np.float32)
        if rowptr is None:
            n, k = ids.shape
            rowptr = np.arange(n + 1, dtype=np.int64) * k
        return np.ascontiguousarray(rowptr, np.int64), ids.reshape(-1), dist.reshape(-1)

    def shortest_paths(self, rowptr, ids, dist, start, end, labels=None):
        rowptr, ids, dist = self._graph_csr(rowptr, ids, dist)
        start = np.ascontiguousarray(start, np.int64)
        end = np.ascontiguousarray(end, np.int64)
        lab = None if labels is None else np.ascontiguousarray(labels, np.int64)
        out = np.zeros(len(start), np.float32)
        self.lib.orc_shortest_paths(len(rowptr) - 1, _p(rowptr, i64p), _p(ids, i64p), _p(dist, f32p), None if lab is None else _p(lab, i64p), len(start),
                                    _p(start, i64p), _p(end, i64p), _p(out, f32p))
        return out

    def sssp(self, rowptr, ids, dist, src):
        rowptr, ids, dist = self._graph_csr(rowptr, ids, dist)
        out = np.zeros(len(rowptr) - 1, np.float32)
        self.lib.orc_sssp(len(rowptr) - 1, _p(rowptr, i64p), _p(ids, i64p), _p(dist, f32p), int(src), _p(out, f32p))
        return out

    def astar(self, ids, dist, x, start, end):
        ids = np.ascontiguousarray(ids, np.int64)
        dist = np.ascontiguousarray(dist, np.float32)
        x = np.ascontiguousarray(x, np.float32)
        start = np.ascontiguousarray(start, np.int64)
        end = np.ascontiguousarray(end, np.int64)
        out = np.zeros(len(start), np.float32)
        self.lib.orc_astar(ids.shape[0], ids.shape[1], _p(ids, i64p), _p(dist, f32p), _p(x, f32p), x.shape[1], len(start), _p(start, i64p), _p(end, i64p),
                           _p(out, f32p))
        return out

    def geodesic_hausdorff(self, rowptr, ids, dist, comp_ptr, comp_pix, pa, pb, labels=None):
        rowptr, ids, dist = self._graph_csr(rowptr, ids, dist)
        comp_ptr = np.ascontiguousarray(comp_ptr, np.int64)
        comp_pix = np.ascontiguousarray(comp_pix, np.int64)
        pa = np.ascontiguousarray(pa, np.int64)
        pb = np.ascontiguousarray(pb, np.int64)
        lab = None if labels is None else np.ascontiguousarray(labels, np.int64)
        out = np.zeros(len(pa), np.float32)
        self.lib.orc_geodesic_hausdorff(len(rowptr) - 1, _p(rowptr, i64p), _p(ids, i64p), _p(dist, f32p), None if lab is None else _p(lab, i64p),
                                        _p(comp_ptr, i64p), _p(comp_pix, i64p), _p(pa, i64p), _p(pb, i64p), len(pa), _p(out, f32p))
        return out

    def prow_graph_ragged(self, rowptr, ids, dist, perplexity=0.0, linear=False):
        rowptr = np.ascontiguousarray(rowptr, np.int64)
        ids = np.ascontiguousarray(ids, np.int64)
        dist = np.ascontiguousarray(dist, np.float32)
        n = len(rowptr) - 1
        m = max(int(rowptr[-1]) - n, 0)
        orp = np.zeros(n + 1, np.int64)
        oc = np.zeros(max(m, 1), np.uint32)
        ov = np.zeros(max(m, 1), np.float32)
        self.lib.orc_prow_graph_ragged(n, _p(rowptr, i64p), _p(ids, i64p), _p(dist, f32p), perplexity, int(linear), _p(orp, i64p), _p(oc, u32p), _p(ov, f32p))
        return orp, oc[:m].copy(), ov[:m].copy()

    def component_centers(self, x, labels, num_components):
        x = np.ascontiguousarray(x, np.float32)
        labels = np.ascontiguousarray(labels, np.int64)
        n, d = x.shape
        centers = np.zeros((num_components, d), np.float32)
        self.lib.orc_component_centers(_p(x, f32p), n, d, _p(labels, i64p), num_components, _p(centers, f32p))
        return centers

    def bridge_components(self, x, labels, num_components, edges, metric=1):
        x = np.ascontiguousarray(x, np.float32)
        labels = np.ascontiguousarray(labels, np.int64)
        edges = np.ascontiguousarray(edges, np.int64).reshape(-1, 2)
        n, d = x.shape
        m = len(edges)
        ia = np.full(m, -1, np.int64)
        ib = np.full(m, -1, np.int64)
        dist = np.zeros(m, np.float32)
        self.lib.orc_bridge_components(_p(x, f32p), n, d, _p(labels, i64p), _p(edges, i64p), m, _p(ia, i64p), _p(ib, i64p), _p(dist, f32p))
        return ia, ib, dist

    def weak_components(self, ids, rowptr=None):
        ids = np.ascontiguousarray(ids, np.int64)
        if rowptr is None:
            n, k = ids.shape
            rowptr = np.arange(n + 1, dtype=np.int64) * k
        rowptr = np.ascontiguousarray(rowptr, np.int64)
        n = len(rowptr) - 1
        labels = np.zeros(max(n, 1), np.int64)
        num = self.lib.orc_weak_components(n, _p(rowptr, i64p), _p(ids.reshape(-1), i64p), _p(labels, i64p))
        return int(num), labels[:n]

    def level_knn_overlap(self, knn_ids, comp_ptr, comp_pix, k):
        knn_ids = np.ascontiguousarray(knn_ids, np.int64)
        comp_ptr = np.ascontiguousarray(comp_ptr, np.int64)
        comp_pix = np.ascontiguousarray(comp_pix, np.int64)
        ncomp = len(comp_ptr) - 1
        ids = np.zeros((ncomp, k), np.int64)
        dist = np.zeros((ncomp, k), np.float32)
        self.lib.orc_level_knn_overlap(_p(knn_ids, i64p), knn_ids.shape[1], ncomp, _p(comp_ptr, i64p), _p(comp_pix, i64p), k, _p(ids, i64p), _p(dist, f32p))
        return ids, dist

    def level_knn_walks(self, rowptr, col, val, k, mode=0):
        rowptr = np.ascontiguousarray(rowptr, np.int64)
        col = np.ascontiguousarray(col, np.uint32)
        val = np.ascontiguousarray(val, np.float32)
        n = len(rowptr) - 1
        ids = np.zeros((n, k), np.int64)
        dist = np.zeros((n, k), np.float32)
        self.lib.orc_level_knn_walks(n, _p(rowptr, i64p), _p(col, u32p), _p(val, f32p), mode, k, _p(ids, i64p), _p(dist, f32p))
        return ids, dist


_oracle = None


def load() -> Oracle:
    global _oracle
    if _oracle is None:
        if not LIB.exists():
            build()
        _oracle = Oracle(C.CDLL(str(LIB)))
    return _oracle


class OracleOps:
    """The compute calls of sph_b200.host.ImageHierarchy served by the CPU oracle (same methods as host.DeviceOps) — lets the
    host merge loop run here without a GPU so that the reference's full-hierarchy goldens pin the oracle."""

    def __init__(self, oracle=None):
        self.o = oracle or load()

    def prow(self, graph, scheme):
        from sph_b200 import host

        n, k = graph.knnIndices.shape
        if scheme == host.NormalizationScheme.TSNE:
            cols, vals, _ = self.o.prow_gaussian_graph(graph.knnIndices, graph.knnDistances, 0.0)
        else:
            cols, vals = self.o.prow_linear_graph(graph.knnIndices, graph.knnDistances)
        return host.SparseRows(np.arange(n + 1, dtype=np.int64) * (k - 1), cols.reshape(-1), vals.reshape(-1))

    def random_walks(self, sims, rws):
        from sph_b200 import host

        rp, c, v = self.o.random_walks(sims.rowptr, sims.col, sims.val, rws.numRandomWalks, rws.singleWalkLength, int(rws.importanceWeighting),
                                       rws.pruneValue, rws.pruneSteps, rws.randomSeed, 1, rws.removeDiagonal, rws.normalize)
        return host.SparseRows(rp, c, v)

    def rw_merge(self, walks, parents, n_merged, norm, weight_by_size):
        from sph_b200 import host

        return host.SparseRows(*self.o.rw_merge(walks.rowptr, walks.col, walks.val, parents, n_merged, norm, weight_by_size))

    def rw_post_merge(self, merged, remove_diagonal, norm):
        from sph_b200 import host

        return host.removeDiagonalAndNormalize(merged, remove_diagonal, norm)

    def pair_sims(self, walks, pa, pb, mode):
        return self.o.pair_sims(walks.rowptr, walks.col, walks.val, pa, pb, mode)

    def merge_graph_nodes(self, graph, parents, n_merged):
        from sph_b200 import host

        return host.RaggedGraph(*self.o.merge_graph_nodes(graph.knnIndices, graph.knnDistances, parents, n_merged, rowptr=graph.rowptr))

    def prow_ragged(self, graph, scheme):
        from sph_b200 import host

        return host.SparseRows(*self.o.prow_graph_ragged(graph.rowptr, graph.knnIndices, graph.knnDistances, 0.0, scheme == host.NormalizationScheme.LINEAR))

    def neighbor_overlap(self, knn_ids, comp_ptr, comp_pix, pa, pb):
        return self.o.neighbor_overlap(knn_ids, comp_ptr, comp_pix, pa, pb)

    def euclid_hausdorff(self, data, comp_ptr, comp_pix, pa, pb):
        return self.o.euclid_hausdorff(data, comp_ptr, comp_pix, pa, pb)

    def geodesic_hausdorff(self, graph, comp_ptr, comp_pix, pa, pb, labels=None):
        return self.o.geodesic_hausdorff(getattr(graph, "rowptr", None), graph.knnIndices, graph.knnDistances, comp_ptr, comp_pix, pa, pb, labels=labels)

    def prow_ragged(self, graph, scheme):
        from sph_b200 import host

        return host.SparseRows(*self.o.prow_graph_ragged(graph.rowptr, graph.knnIndices, graph.knnDistances, 0.0, scheme == host.NormalizationScheme.LINEAR))


def dense_to_csr(m):
    m = np.asarray(m, np.float32)
    rowptr = [0]
    col, val = [], []
    for r in m:
        nz = np.nonzero(r)[0]
        col.extend(nz.tolist())
        val.extend(r[nz].tolist())
        rowptr.append(len(col))
    return np.array(rowptr, np.int64), np.array(col, np.uint32), np.array(val, np.float32)


def csr_to_dense(rowptr, col, val, ncols=None):
    n = len(rowptr) - 1
    ncols = n if ncols is None else ncols
    out = np.zeros((n, ncols), np.float32)
    for i in range(n):
        for e in range(rowptr[i], rowptr[i + 1]):
            out[i, col[e]] = val[e]
    return out
