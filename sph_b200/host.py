"""Host-side mirror of the reference's stage interface for the hot path, above the C ABI.

The reference is C++ (`sph::NearestNeighbors`, `sph::LevelSimilarities`, the free functions in
`sph::utils`); its own toolchain dependencies (faiss, Eigen, HDILib, Boost) are absent here, so the
C++ shim a maintainer would add is shown in INTEGRATION.md and this module is the Python restatement
of the same call sequence — same names, same argument meaning, same error behaviour (log + return,
outputs untouched) — used by the tests and bench.py.  Every compute call goes through
`sph_b200.capi` (ctypes over include/sphb.h); there is no CPU path in here.

Multi-GPU: query rows are independent, so `shard_rows` splits them into contiguous slabs,
every rank runs the same kNN + P-row call on its slab against the replicated database, and
`gather_rows` concatenates the slabs (one all_gather per output, no data-path collective inside the
kernels).
"""
from __future__ import annotations

import enum
import logging
from dataclasses import dataclass, field

import numpy as np

log = logging.getLogger("sph_b200")


# ---- sph/utils/Settings.hpp:34-181 ------------------------------------------------------------
class KnnIndex(enum.IntEnum):
    BruteForce = 0
    Flat = 1
    IVFFlat = 2
    HNSW = 3
    HNSWSQ = 4
    HNSW_IVFPQ = 5


class KnnMetric(enum.IntEnum):
    L2 = 0
    COSINE = 1
    INNER_PRODUCT = 2


class ImportanceWeighting(enum.IntEnum):
    CONSTANT = 0
    LINEAR = 1
    NORMAL = 2
    ONLYLAST = 3
    FIRST_VISIT = 4


class NormalizationScheme(enum.IntEnum):
    NONE = 0
    TSNE = 1
    UMAP = 2
    LINEAR = 3


@dataclass
class NearestNeighborsSettings:  # Settings.hpp:134-143
    numNearestNeighbors: int = 0
    knnIndex: KnnIndex = KnnIndex.Flat
    knnMetric: KnnMetric = KnnMetric.L2
    symmetricNeighbors: bool = False
    computeConnectComponents: bool = False
    neighborConnectComponents: bool = False
    L2squared: bool = False


@dataclass
class RandomWalkSettings:  # Settings.hpp:91-103
    numRandomWalks: int = 90
    singleWalkLength: int = 15
    minimumSingleWalkLength: int = 5
    pruneValue: float = 0.0
    pruneSteps: int = 0
    importanceWeighting: ImportanceWeighting = ImportanceWeighting.CONSTANT
    normalize: bool = True
    removeDiagonal: bool = True
    randomSeed: int = 1
    parallel: bool = True


@dataclass
class LevelSimilaritiesSettings:  # Settings.hpp:169-181 (the fields this path reads)
    ks: list = field(default_factory=list)
    exactKnn: bool = True
    levelToCompute: int = -1
    randomWalkPairSims: bool = True
    weightTransitionBySize: bool = False
    normalizeProbDist: NormalizationScheme = NormalizationScheme.TSNE
    computeSymmetricProbDist: NormalizationScheme = NormalizationScheme.TSNE


@dataclass
class Graph:  # sph/utils/Graph.hpp:155-273 with a fixed number of neighbours
    knnIndices: np.ndarray
    knnDistances: np.ndarray
    symmetric: bool = False

    @property
    def numPoints(self) -> int:
        return self.knnIndices.shape[0]

    @property
    def nns(self) -> np.ndarray:
        return np.full(self.numPoints, self.knnIndices.shape[1], np.int64)

    @property
    def offsets(self) -> np.ndarray:  # exclusive scan (Graph.hpp:254-258)
        return np.arange(self.numPoints, dtype=np.int64) * self.knnIndices.shape[1]

    def isValid(self) -> bool:
        return self.knnIndices.shape == self.knnDistances.shape and bool((self.knnIndices >= 0).all())


@dataclass
class SparseRows:
    """Flat form of SparseMatHDI / SparseMatSPH (sph/utils/CommonDefinitions.hpp:87-93): CSR with column-sorted rows."""
    rowptr: np.ndarray
    col: np.ndarray
    val: np.ndarray

    @property
    def n(self) -> int:
        return len(self.rowptr) - 1

    def row(self, i: int):
        b, e = int(self.rowptr[i]), int(self.rowptr[i + 1])
        return self.col[b:e], self.val[b:e]


# ---- sharding -----------------------------------------------------------------------------------
def shard_rows(n: int, world: int) -> list[tuple[int, int]]:
    """Contiguous query-row slabs, sizes differing by at most one (rank r gets [n*r//world, n*(r+1)//world))."""
    if world <= 0:
        raise ValueError("world must be positive")
    return [((n * r) // world, (n * (r + 1)) // world) for r in range(world)]


def gather_rows(local: np.ndarray, n: int, group=None) -> np.ndarray:
    """Concatenates the ranks' row slabs (shapes [nq_r, ...]) into [n, ...] on every rank."""
    import torch
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return local
    world = dist.get_world_size(group)
    slabs = shard_rows(n, world)
    dev = "cuda" if dist.get_backend(group) == "nccl" else "cpu"
    maxq = max(b - a for a, b in slabs)
    buf = torch.zeros((maxq,) + tuple(local.shape[1:]), dtype=torch.from_numpy(local[:0]).dtype, device=dev)
    buf[: local.shape[0]] = torch.from_numpy(np.ascontiguousarray(local)).to(dev)
    parts = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(parts, buf, group=group)
    return np.concatenate([p[: b - a].cpu().numpy() for p, (a, b) in zip(parts, slabs)], axis=0)


# ---- NearestNeighbors (sph/NearestNeighbors.hpp:32-125, .cpp:98-189) ----------------------------
class NearestNeighbors:
    """Data-level exact kNN graph.  `compute()` follows NearestNeighbors::compute: metric setup (:81-96),
    COSINE renormalisation (:208-222), exact search, post-passes (:152-167, fused into the CUDA write-out)."""

    def __init__(self, data: np.ndarray, knn_fn=None, prow_fn=None):
        self._data = np.ascontiguousarray(data, np.float32)
        self._graph: Graph | None = None
        self._nns = NearestNeighborsSettings()
        self._valid = False
        self._knn_fn = knn_fn  # injectable for the CPU multi-process tests (the product default is the C ABI)

    def compute(self, nns: NearestNeighborsSettings | None = None, rank: int = 0, world: int = 1, group=None) -> None:
        if nns is not None:
            self._nns = nns
        s = self._nns
        n, _ = self._data.shape
        self._valid = False
        if s.numNearestNeighbors <= 0:
            log.error("NearestNeighbors::compute: numNearestNeighbors must be positive")
            return
        if s.knnIndex not in (KnnIndex.BruteForce, KnnIndex.Flat):
            # the approximate indices are replaced by the exact GPU search
            log.info("NearestNeighbors::compute: index %s served by the exact GPU search", s.knnIndex.name)
        x = self._data
        metric, l2sq = 1, s.L2squared  # faiss::METRIC_L2
        if s.knnMetric in (KnnMetric.COSINE, KnnMetric.INNER_PRODUCT):
            metric, l2sq = 0, False  # setMetric: IP / COSINE force L2squared = false (:81-96)
            if s.knnMetric == KnnMetric.COSINE:
                x = renorm_l2(x)
        q0, q1 = shard_rows(n, world)[rank]
        try:
            ids, dist = self._knn(x, int(s.numNearestNeighbors), metric, l2sq, q0, q1)
        except Exception as e:  # reference convention: log + early return, outputs untouched
            log.error("NearestNeighbors::compute failed: %s", e)
            return
        if world > 1:
            ids = gather_rows(ids, n, group)
            dist = gather_rows(dist, n, group)
        self._graph = Graph(ids, dist)
        self._valid = checkAllNeighborsExist(ids) or n < s.numNearestNeighbors

    def _knn(self, x, k, metric, l2sq, q0, q1):
        if self._knn_fn is not None:
            return self._knn_fn(x, k, metric, l2sq, q0, q1)
        from . import capi

        return capi.knn_exact(x, k, metric=metric, l2_squared=l2sq, q0=q0, q1=q1)

    def isValid(self) -> bool:
        return self._valid

    def getKnnGraph(self) -> Graph:
        return self._graph

    def getKnnIndices(self) -> np.ndarray:
        return self._graph.knnIndices

    def getKnnDistances(self) -> np.ndarray:
        return self._graph.knnDistances

    @staticmethod
    def indexHeuristic(n: int) -> KnnIndex:  # sph/NearestNeighbors.hpp:50-63 (kept for API parity; every choice runs exact here)
        if n <= 10_000:
            return KnnIndex.BruteForce
        if n <= 100_000:
            return KnnIndex.IVFFlat
        if n <= 25_000_000:
            return KnnIndex.HNSW
        return KnnIndex.HNSW_IVFPQ


def renorm_l2(x: np.ndarray) -> np.ndarray:
    """faiss::fvec_renorm_L2 on a copy (sph/NearestNeighbors.cpp:208-222): fp32 sum of squares, scale by 1/sqrt."""
    out = np.array(x, np.float32, copy=True)
    nr = np.zeros(len(out), np.float32)
    for c in range(out.shape[1]):  # sequential fp32 accumulation, column by column
        nr += out[:, c] * out[:, c]
    pos = nr > 0
    out[pos] *= (np.float32(1.0) / np.sqrt(nr[pos]))[:, None]
    return out


def checkAllNeighborsExist(ids: np.ndarray) -> bool:  # sph/utils/Knn.cpp:370-398
    return bool((ids != -1).all())


def numNeighborsFromSetting(n: int, nKnns: int) -> int:
    """evaluation/RunEvaluation.cpp:241-248: nKnns == 0 -> clamp(N/100, 10, 100) * 3; then +1 for the point itself."""
    k = nKnns if nKnns > 0 else int(min(max(n // 100, 10), 100)) * 3
    return k + 1


# ---- P rows (sph/utils/GraphNormalization.hpp:16-53) ---------------------------------------------
def normalizeKnnDistances(graph: Graph, scheme: NormalizationScheme, perplexity: float = 0.0) -> SparseRows:
    """utils::normalizeKnnDistances dispatcher: TSNE -> Gaussian rows, LINEAR -> linear rows."""
    from . import capi

    n, k = graph.knnIndices.shape
    if scheme == NormalizationScheme.TSNE:
        cols, vals = capi.prow_gaussian_graph(graph.knnIndices, graph.knnDistances, perplexity)
    elif scheme == NormalizationScheme.LINEAR:
        cols, vals = capi.prow_linear_graph(graph.knnIndices, graph.knnDistances)
    else:
        raise NotImplementedError(f"normalizeKnnDistances: {scheme.name} is a §8(f) 'next' row")
    rowptr = np.arange(n + 1, dtype=np.int64) * (k - 1)
    return SparseRows(rowptr, cols.reshape(-1), vals.reshape(-1))


def level_schedule(level_sizes, k0: int):
    """LevelSimilarities::updateNumberOfNeighbors (sph/LevelSimilarities.cpp:83-115): (perplexities, ks)."""
    f = np.float32
    base = f(k0 - 1) / f(3.0)
    perps = [float(min(max(base, f(10.0)), f(100.0)))]
    ks = [int(k0)]
    for n_l in list(level_sizes)[1:]:
        p = min(max(f(n_l) / f(100.0), f(10.0)), f(100.0))
        p = min(base, p)
        perps.append(float(p))
        ks.append(int(min(int(p) * 3 + 1, n_l)))
    return perps, ks


# ---- LevelSimilarities (sph/LevelSimilarities.hpp:47-97, .cpp:444-623) ---------------------------
class LevelSimilarities:
    """Probability rows per level.  Level 0 comes from the kNN graph (computeGaussianDistributions, :531);
    level >= 1 from the merged random walks (useRandomWalks, :460-508): createSimilarities (bottom-k of the
    Bhattacharyya distances) or the top-k transition values, then the in-place Gaussian rows with ignore = -1,
    zero removal (:566-581) and symmetrizeOutput (:589-623)."""

    def __init__(self, knn_graph: Graph, random_walks: list[SparseRows] | None = None):
        self._graph = knn_graph
        self._rw = random_walks or []
        self._prob: dict[int, SparseRows] = {}
        self._perplexities: list[float] = []
        self._ks: list[int] = []

    def compute(self, lss: LevelSimilaritiesSettings | None = None) -> None:
        from . import capi

        lss = lss or LevelSimilaritiesSettings()
        sizes = [self._graph.numPoints] + [rw.n for rw in self._rw[1:]]
        self._perplexities, self._ks = level_schedule(sizes, self._graph.knnIndices.shape[1])
        levels = range(len(sizes)) if lss.levelToCompute < 0 else [lss.levelToCompute]
        for level in levels:
            if level >= len(sizes):
                log.error("LevelSimilarities::compute: level %d does not exist", level)
                return
            if level == 0:
                self._prob[0] = normalizeKnnDistances(self._graph, NormalizationScheme.TSNE, self._perplexities[0])
                continue
            rw = self._rw[level]
            k_l, perp_l = self._ks[level], self._perplexities[level]
            if lss.randomWalkPairSims:
                cols, vals, nnz = capi.create_similarities(rw.rowptr, rw.col, rw.val, k_l, 1e-4)
            else:
                raise NotImplementedError("randomWalkPairSims = false (top-k transition values) is not wired yet")
            rowptr = np.concatenate([[0], np.cumsum(nnz, dtype=np.int64)])
            keep = np.arange(cols.shape[1])[None, :] < nnz[:, None]
            c, v = cols[keep], vals[keep]
            if lss.normalizeProbDist == NormalizationScheme.TSNE:
                v = capi.prow_gaussian_inplace(rowptr, v, perp_l, -1)
            nz = v != 0  # erase exact zeros (:566-581)
            counts = np.add.reduceat(nz.astype(np.int64), rowptr[:-1]) if len(v) else np.zeros(len(rowptr) - 1, np.int64)
            counts[np.diff(rowptr) == 0] = 0
            self._prob[level] = SparseRows(np.concatenate([[0], np.cumsum(counts)]).astype(np.int64), c[nz], v[nz])
        self.symmetrizeOutput(lss.computeSymmetricProbDist)

    def symmetrizeOutput(self, scheme: NormalizationScheme) -> None:
        from . import capi

        if scheme not in (NormalizationScheme.TSNE, NormalizationScheme.UMAP):
            return
        for level, p in list(self._prob.items()):
            rp, c, v = capi.symmetrize(p.rowptr, p.col, p.val, umap=scheme == NormalizationScheme.UMAP)
            self._prob[level] = SparseRows(rp, c, v)

    def getProbDist(self, level: int) -> SparseRows:
        return self._prob[level]

    def getPerplexities(self):
        return self._perplexities

    def getKs(self):
        return self._ks


# ---- random walks (sph/utils/SparseMatrixAlgorithms.hpp:73,82) -----------------------------------
def doRandomWalks(similarities: SparseRows, rws: RandomWalkSettings) -> SparseRows:
    from . import capi

    rp, c, v = capi.random_walks(similarities.rowptr, similarities.col, similarities.val, rws.numRandomWalks, rws.singleWalkLength,
                                 int(rws.importanceWeighting), rws.pruneValue, rws.pruneSteps, rws.randomSeed, rws.removeDiagonal, rws.normalize)
    return SparseRows(rp, c, v)


def mergeNodesRandomWalks(walks: SparseRows, num_merged: int, parents, norm: bool = True, weight_by_size: bool = True) -> SparseRows:
    from . import capi

    rp, c, v = capi.rw_merge(walks.rowptr, walks.col, walks.val, parents, num_merged, norm, weight_by_size)
    return SparseRows(rp, c, v)
