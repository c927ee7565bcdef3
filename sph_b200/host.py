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
    componentSim: "ComponentSim | None" = None   # None: NEIGH_WALKS with MERGE_RW_ONLY (useRandomWalks)
    knnFromWalkMetric: bool = False               # rwHandling == MERGE_RW_NEW_WALKS_AND_KNN: exact level kNN + useKnnDistances
    forceComputeDistances: bool = False           # useRandomWalksAsKnnDistances
    componentLabels: np.ndarray | None = None     # GEO_CENTROID: connected-component label per point (sph/ComputeHierarchy.cpp:144)


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
class RaggedGraph:  # sph/utils/Graph.hpp:155-273 with per-point neighbour counts (symmetrised / merged / level graphs)
    rowptr: np.ndarray       # offsets + total
    knnIndices: np.ndarray   # int64 [nnz]
    knnDistances: np.ndarray  # float32 [nnz]
    symmetric: bool = False

    @property
    def numPoints(self) -> int:
        return len(self.rowptr) - 1

    @property
    def nns(self) -> np.ndarray:
        return np.diff(self.rowptr)

    def getNeighbors(self, p: int) -> np.ndarray:
        return self.knnIndices[self.rowptr[p]:self.rowptr[p + 1]]

    def getDistances(self, p: int) -> np.ndarray:
        return self.knnDistances[self.rowptr[p]:self.rowptr[p + 1]]


class ComponentSim(enum.IntEnum):  # sph/utils/Settings.hpp (ComponentSim)
    NEIGH_OVERLAP = 0
    GEO_CENTROID = 1
    NEIGH_WALKS = 2
    NEIGH_WALKS_SINGLE_OVERLAP = 3
    EUCLID_CENTROID = 4
    GEO_WALKS = 5


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


# ---- host parts of NearestNeighbors::connectComponents (sph/NearestNeighbors.cpp:494-861) ----------------------------
def centroid_spanning_tree(centers: np.ndarray) -> np.ndarray:
    """computeSpanningTree (:682-708): minimum spanning tree of the complete graph over the component centres, edge weight
    utils::L2 (sqrt of the fp32 squared distance, sph/utils/Distances.hpp:40-43).  Returns int64 [ncomp-1, 2]: (source, target)
    with source < target — Boost stores the edge (i, j) as added in the i < j loop (:690-694) — in ascending order of weight,
    the order kruskal_minimum_spanning_tree emits them.  With pairwise distinct weights the tree is unique; among exactly equal
    weights Boost's priority queue order is unspecified, here ties go to the lower (source, target).  Prim over on-the-fly
    distance rows: O(ncomp^2 d) time, O(ncomp) memory (the reference materialises all ncomp^2 / 2 Boost edges)."""
    c = np.ascontiguousarray(centers, np.float32)
    m = len(c)
    if m <= 1:
        return np.zeros((0, 2), np.int64)
    in_tree = np.zeros(m, bool)
    best = np.full(m, np.inf, np.float32)
    link = np.full(m, -1, np.int64)
    out = []
    v = 0
    for _ in range(m - 1):
        in_tree[v] = True
        diff = c - c[v]
        w = np.sqrt(np.einsum("ij,ij->i", diff, diff, dtype=np.float32))
        upd = ~in_tree & ((w < best) | ((w == best) & (v < link)))
        best[upd] = w[upd]
        link[upd] = v
        cand = np.where(in_tree, np.inf, best)
        v = int(np.argmin(cand))
        a, b = (int(link[v]), v) if link[v] < v else (v, int(link[v]))
        out.append((float(best[v]), a, b))
    out.sort()
    return np.array([(a, b) for _, a, b in out], np.int64).reshape(-1, 2)


def insert_graph_edges(rowptr, ids, dist, ids_a, ids_b, dd):
    """insertDistance (:547-572) for (a, b, d) and (b, a, d) of every bridge, in the given order: skipped when a == b or b is
    already a neighbour of a; otherwise inserted behind the last distance <= d (std::upper_bound), never at position 0."""
    touched: dict[int, tuple[list, list]] = {}

    def row(a):
        if a not in touched:
            b0, b1 = int(rowptr[a]), int(rowptr[a + 1])
            touched[a] = (ids[b0:b1].tolist(), dist[b0:b1].tolist())
        return touched[a]

    def insert(a, b, d):
        if a == b or a < 0 or b < 0 or a >= len(rowptr) - 1:
            return
        rid, rd = row(a)
        if b in rid:
            return
        pos = int(np.searchsorted(np.asarray(rd, np.float32), np.float32(d), side="right"))
        pos = max(pos, 1)
        rid.insert(pos, b)
        rd.insert(pos, float(np.float32(d)))

    for a, b, d in zip(ids_a.tolist(), ids_b.tolist(), dd.tolist()):
        insert(a, b, d)
        insert(b, a, d)
    if not touched:
        return rowptr, ids, dist
    n = len(rowptr) - 1
    nns = np.diff(rowptr)
    for a, (rid, _) in touched.items():
        nns[a] = len(rid)
    new_rowptr = np.concatenate([[0], np.cumsum(nns)]).astype(np.int64)
    new_ids = np.empty(new_rowptr[-1], np.int64)
    new_dist = np.empty(new_rowptr[-1], np.float32)
    keep = np.ones(n, bool)
    keep[list(touched)] = False
    # untouched rows move as blocks: every old entry shifts by the growth of the rows before it
    shift = np.repeat(new_rowptr[:-1] - rowptr[:-1], np.diff(rowptr))
    mask = np.repeat(keep, np.diff(rowptr))
    dst = (np.arange(len(ids), dtype=np.int64) + shift)[mask]
    new_ids[dst] = ids[mask]
    new_dist[dst] = dist[mask]
    for a, (rid, rd) in touched.items():
        b0 = int(new_rowptr[a])
        new_ids[b0:b0 + len(rid)] = rid
        new_dist[b0:b0 + len(rd)] = rd
    return new_rowptr, new_ids, new_dist


def ensureClosestPointIsSelf(g) -> int:
    """utils::ensureClosestPointIsSelf (sph/utils/GraphUtils.cpp:23-95) on a RaggedGraph, in place; returns the number of adjusted rows.
    (Stage A and symmetrizeGraph apply it on the device; this host form serves the connected graph, where it is a no-op unless an
    upstream row was already broken.)"""
    first = g.knnIndices[g.rowptr[:-1]] if g.numPoints else np.zeros(0, np.int64)
    bad = np.nonzero((first != np.arange(g.numPoints)) & (np.diff(g.rowptr) > 0))[0]
    for p in bad.tolist():
        b0, b1 = int(g.rowptr[p]), int(g.rowptr[p + 1])
        nb, ds = g.knnIndices[b0:b1], g.knnDistances[b0:b1]
        hit = np.nonzero(nb == p)[0]
        if hit.size:
            sp = int(hit[0])
        elif (ds == 0).all():
            sp = len(nb) - 1
            nb[sp] = p
        else:
            nb[1:] = nb[:-1].copy()
            ds[1:] = ds[:-1].copy()
            sp = 0
            nb[0], ds[0] = p, 0.0
        nb[0], nb[sp] = nb[sp], nb[0]
        ds[0], ds[sp] = ds[sp], ds[0]
    return int(bad.size)


# ---- NearestNeighbors (sph/NearestNeighbors.hpp:32-125, .cpp:98-189) ----------------------------
class NearestNeighbors:
    """Data-level exact kNN graph.  `compute()` follows NearestNeighbors::compute: metric setup (:81-96),
    COSINE renormalisation (:208-222), exact search, post-passes (:152-167, fused into the CUDA write-out)."""

    def __init__(self, data: np.ndarray, knn_fn=None, prow_fn=None):
        self._data = np.ascontiguousarray(data, np.float32)
        self._graph: Graph | None = None
        self._sym_graph: RaggedGraph | None = None
        self._cc = None
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
                metric = 2  # SPHB_METRIC_COSINE: the back-end normalises a device copy of the rows (initFaissData) and searches by inner product
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
        self._sym_graph = None
        if s.symmetricNeighbors and self._valid:
            self.computeSymmetrizedNnGraph()
        self._cc = None
        if s.computeConnectComponents and self._valid:
            self.computeConnectedComponents()

    def computeSymmetrizedNnGraph(self):
        """NearestNeighbors::computeSymmetrizedNnGraph (sph/NearestNeighbors.cpp:411-492) -> utils::symmetrizeGraph."""
        if self._graph is None or self._graph.numPoints == 0:
            log.warning("NearestNeighbors::computeSymmetrizedNnGraph: _knn graph is not computed yet, doing nothing")
            return None
        if not checkAllNeighborsExist(self._graph.knnIndices):
            log.error("NearestNeighbors::computeSymmetrizedNnGraph: missing neighbours")
            return None
        try:
            rp, ids, dist = self._symmetrize(self._graph.knnIndices, self._graph.knnDistances)
        except Exception as e:
            log.error("NearestNeighbors::computeSymmetrizedNnGraph failed: %s", e)
            return None
        self._sym_graph = RaggedGraph(rp, ids, dist, symmetric=True)
        return self._sym_graph

    def computeConnectedComponents(self):
        """NearestNeighbors::computeConnectedComponents (sph/NearestNeighbors.cpp:318-409): strong components of the symmetric
        graph when it exists, else weak components of the kNN graph — both are the undirected components."""
        from . import capi

        if self._graph is None:
            log.warning("NearestNeighbors::computeConnectedComponents: _knn graph is not computed yet, call compute() first")
            return None
        try:
            if self._sym_graph is not None:
                self._cc = capi.weak_components(self._sym_graph.knnIndices, self._sym_graph.rowptr)
            else:
                self._cc = capi.weak_components(self._graph.knnIndices)
        except Exception as e:
            log.error("NearestNeighbors::computeConnectedComponents failed: %s", e)
            return None
        return self._cc

    def getConnectedComponents(self):
        return self._cc

    def connectComponents(self, bridge_ops=None):
        """NearestNeighbors::connectComponents (sph/NearestNeighbors.cpp:494-861): one extra bidirectional edge per edge of the
        minimum spanning tree over the component centres, between the closest pair of points of the two components.
        Device: centres of mass and the closest pairs of ALL tree edges (sphb_component_centers, sphb_bridge_components);
        host: the tree itself (:682-708) and the row insertions (:547-572), as SURVEY §8(f)-2 splits it.
        bridge_ops: object with component_centers / bridge_components (tests inject the CPU oracle); default the C ABI."""
        if self._graph is None:
            log.warning("NearestNeighbors::connectComponents: _knn graph is not computed yet, call compute() first")
            return None
        if self._cc is None and self.computeConnectedComponents() is None:
            return None
        if bridge_ops is None:
            from . import capi as bridge_ops
        num, labels = self._cc
        if self._sym_graph is not None:
            g = self._sym_graph
            rowptr, ids, dist, sym = g.rowptr.copy(), g.knnIndices.copy(), g.knnDistances.copy(), True
        else:
            n, k = self._graph.knnIndices.shape
            rowptr = np.arange(n + 1, dtype=np.int64) * k
            ids, dist, sym = self._graph.knnIndices.reshape(-1).copy(), self._graph.knnDistances.reshape(-1).copy(), False
        if num > 1:
            if self._nns.knnMetric != KnnMetric.L2:
                log.error("NearestNeighbors::connectComponents: only the L2 metric is served by the device search")
                return None
            try:
                centers = bridge_ops.component_centers(self._data, labels, num)
                edges = centroid_spanning_tree(centers)
                ia, ib, dd = bridge_ops.bridge_components(self._data, labels, num, edges)
            except Exception as e:
                log.error("NearestNeighbors::connectComponents failed: %s", e)
                return None
            rowptr, ids, dist = insert_graph_edges(rowptr, ids, dist, ia, ib, dd)
        self._connected_graph = RaggedGraph(rowptr, ids, dist, symmetric=sym)
        ensureClosestPointIsSelf(self._connected_graph)
        return self._connected_graph

    def getConnectedGraph(self):
        return getattr(self, "_connected_graph", None)

    def _symmetrize(self, ids, dist):
        from . import capi

        return capi.symmetrize_graph(ids, dist)

    def getSymGraph(self):
        return self._sym_graph

    def _knn(self, x, k, metric, l2sq, q0, q1):
        if self._knn_fn is not None:
            if metric == 2:  # an injected search (tests: the CPU oracle) gets the normalised copy + inner product, as faiss does
                x, metric = renorm_l2(x), 0
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

    def __init__(self, knn_graph: Graph, random_walks: list[SparseRows] | None = None, hierarchy=None):
        self._graph = knn_graph
        self._rw = random_walks or (hierarchy.randomWalks if hierarchy is not None else [])
        self._hierarchy = hierarchy  # an ImageHierarchy: component -> pixel maps for the neighbour-overlap metric
        self._prob: dict[int, SparseRows] = {}
        self._perplexities: list[float] = []
        self._ks: list[int] = []
        self._distance_graphs: dict[int, object] = {}

    def computeNearestNeighborOnLevel(self, level: int, lss: LevelSimilaritiesSettings):
        """LevelSimilarities::computeNearestNeighborOnLevel (sph/LevelSimilarities.cpp:184-442) for the random-walk metrics:
        exact kNN in the component metric (computeExactKnn, :211-252) or the walks themselves as distances (:346-389)."""
        from . import capi

        if lss.componentSim == ComponentSim.NEIGH_OVERLAP:
            ptr, pix = self._hierarchy.representedDataPoints(level)
            k_l = min(self._ks[level], len(ptr) - 1)
            ids, dist = capi.level_knn_overlap(self._graph.knnIndices, ptr, pix, k_l)
            self._distance_graphs[level] = Graph(ids, dist)
            return self._distance_graphs[level]
        if lss.componentSim == ComponentSim.EUCLID_CENTROID:
            # EuclidDistSpace (sph/utils/EuclidDistSpace.hpp:29-39): Hausdorff distance of every component pair, the k smallest
            # (distance, index) per row selected on the device (sphb_level_knn_euclid)
            ptr, pix = self._hierarchy.representedDataPoints(level)
            n_l = len(ptr) - 1
            k_l = min(self._ks[level], n_l)
            ids, dist = capi.level_knn_euclid(self._hierarchy._data, ptr, pix, k_l)
            self._distance_graphs[level] = Graph(ids, dist)
            return self._distance_graphs[level]
        if lss.componentSim == ComponentSim.GEO_CENTROID:
            # GeodesicPathSpace (sph/utils/GeodesicPathSpace.hpp:38-49): geodesicDistance of every component pair — one batched
            # device call — then the k smallest (distance, index) per row
            ptr, pix = self._hierarchy.representedDataPoints(level)
            n_l = len(ptr) - 1
            k_l = min(self._ks[level], n_l)
            pa = np.repeat(np.arange(n_l, dtype=np.int64), n_l)
            pb = np.tile(np.arange(n_l, dtype=np.int64), n_l)
            sptr, spix, spa, spb = subsampleComponentPairs(ptr, pix, pa, pb, self._hierarchy.ihs.numGeodesicSamples)
            dm = geodesicDistance(self._graph, sptr, spix, spa, spb, lss.componentLabels).reshape(n_l, n_l)
            order = np.lexsort((pb.reshape(n_l, n_l), dm), axis=1)[:, :k_l]
            self._distance_graphs[level] = Graph(order.astype(np.int64), np.take_along_axis(dm, order, 1).astype(np.float32))
            return self._distance_graphs[level]
        rw = self._rw[level]
        k_l = min(self._ks[level], rw.n)
        if lss.knnFromWalkMetric:
            mode = 1 if lss.componentSim == ComponentSim.NEIGH_WALKS_SINGLE_OVERLAP else 0
            ids, dist = capi.level_knn_walks(rw.rowptr, rw.col, rw.val, k_l, mode)
            self._distance_graphs[level] = Graph(ids, dist)
        elif lss.forceComputeDistances:
            rp, ids, dist = capi.rw_as_knn_distances(rw.rowptr, rw.col, rw.val)
            self._distance_graphs[level] = RaggedGraph(rp, ids, dist)
        return self._distance_graphs.get(level)

    def getSimilaritiesGraph(self, level: int):
        return self._distance_graphs.get(level)

    def compute(self, lss: LevelSimilaritiesSettings | None = None, rank: int = 0, world: int = 1, group=None) -> None:
        """LevelSimilarities::compute (sph/LevelSimilarities.cpp:118-182).  world > 1 (one process per GPU): createSimilarities is
        sharded by output row over the ranks (createSimilarities above); every rank ends with the whole level."""
        from . import capi

        lss = lss or LevelSimilaritiesSettings()
        if self._hierarchy is not None:
            sizes = list(self._hierarchy.numComponents)
        else:
            sizes = [self._graph.numPoints] + [rw.n for rw in self._rw[1:]]
        if lss.ks:  # _lss.ks[0] is the number of data-level neighbours (LevelSimilarities.cpp:96-98)
            k0 = int(lss.ks[0])
        elif self._graph.knnIndices.ndim == 2:
            k0 = self._graph.knnIndices.shape[1]
        else:
            log.error("LevelSimilarities::compute: a ragged data graph needs lss.ks[0], the number of data-level neighbours")
            return
        self._perplexities, self._ks = level_schedule(sizes, k0)
        levels = range(len(sizes)) if lss.levelToCompute < 0 else [lss.levelToCompute]
        for level in levels:
            if level >= len(sizes):
                log.error("LevelSimilarities::compute: level %d does not exist", level)
                return
            if level == 0:
                if isinstance(self._graph, RaggedGraph):
                    g = self._graph
                    self._prob[0] = SparseRows(*capi.prow_graph_ragged(g.rowptr, g.knnIndices, g.knnDistances, self._perplexities[0], False))
                else:
                    self._prob[0] = normalizeKnnDistances(self._graph, NormalizationScheme.TSNE, self._perplexities[0])
                continue
            overlap = lss.componentSim in (ComponentSim.NEIGH_OVERLAP, ComponentSim.EUCLID_CENTROID, ComponentSim.GEO_CENTROID)  # computeKnn + useKnnDistances
            if overlap and self._hierarchy is None:
                log.error("LevelSimilarities::compute: %s needs the image hierarchy", lss.componentSim.name)
                return
            rw = None if overlap else self._rw[level]
            k_l, perp_l = min(self._ks[level], sizes[level]), self._perplexities[level]
            if overlap or lss.knnFromWalkMetric or lss.forceComputeDistances:
                self.computeNearestNeighborOnLevel(level, lss)
            if overlap or lss.knnFromWalkMetric:
                # useKnnDistances (:510-515): flat Gaussian rows (tol 1e-5, ignore 0) scattered by neighbour id, skipping -1
                g = self._distance_graphs[level]
                temp = capi.prow_gaussian_flat(g.knnDistances, perp_l)
                rowptr, c, v = scatter_flat_rows(g.knnIndices, temp)
            else:
                if lss.randomWalkPairSims:
                    cols, vals, nnz = createSimilarities(rw, k_l, 1e-4, rank=rank, world=world, group=group)
                else:  # convertEigenSparseVecToHDILibSparseVec: the k_l largest transition values per row (:481-486)
                    cols, vals, nnz = capi.find_k(rw.rowptr, rw.col, rw.val, k_l, top=True)
                rowptr = np.concatenate([[0], np.cumsum(nnz, dtype=np.int64)])
                keep = np.arange(cols.shape[1])[None, :] < nnz[:, None]
                c, v = cols[keep], vals[keep]
                if lss.normalizeProbDist == NormalizationScheme.TSNE:
                    v = capi.prow_gaussian_inplace(rowptr, v, perp_l, -1)
            rp, c, v = capi.remove_zeros(rowptr, c, v)  # erase exact zeros (:566-581)
            self._prob[level] = SparseRows(rp, c, v)
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


def scatter_flat_rows(ids: np.ndarray, temp: np.ndarray):
    """The serial scatter of utils::computeGaussianDistributions (sph/utils/HDILibHelper.hpp:131-139): slots 1..k-1 of every
    row go to `distribution[j][ids]` (a column-sorted map; a later duplicate id overwrites), ids == -1 are skipped."""
    n, k = ids.shape
    rowptr = [0]
    cols, vals = [], []
    for j in range(n):
        row = {}
        for t in range(1, k):
            if ids[j, t] != -1:
                row[int(ids[j, t])] = temp[j, t]
        for cc in sorted(row):
            cols.append(cc)
            vals.append(row[cc])
        rowptr.append(len(cols))
    return np.array(rowptr, np.int64), np.array(cols, np.uint32), np.array(vals, np.float32)


# ---- random walks (sph/utils/SparseMatrixAlgorithms.hpp:73,82) -----------------------------------
def gather_csr(local: SparseRows, n: int, group=None) -> SparseRows:
    """Concatenates the ranks' CSR row slabs (rank r holds rows shard_rows(n, world)[r], rowptr relative to its slab)."""
    import torch
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return local
    counts = gather_rows(np.diff(local.rowptr).astype(np.int64), n, group)
    world = dist.get_world_size(group)
    dev = "cuda" if dist.get_backend(group) == "nccl" else "cpu"
    slabs = shard_rows(n, world)
    sizes = [int(counts[a:b].sum()) for a, b in slabs]
    mx = max(max(sizes), 1)

    def gather_flat(x, dtype):
        buf = torch.zeros(mx, dtype=dtype, device=dev)
        buf[: len(x)] = torch.from_numpy(np.ascontiguousarray(x).astype(np.int64 if dtype == torch.int64 else np.float32)).to(dev)
        parts = [torch.empty_like(buf) for _ in range(world)]
        dist.all_gather(parts, buf, group=group)
        return np.concatenate([p[:sz].cpu().numpy() for p, sz in zip(parts, sizes)])

    col = gather_flat(local.col, torch.int64).astype(np.uint32)
    val = gather_flat(local.val, torch.float32).astype(np.float32)
    return SparseRows(np.concatenate([[0], np.cumsum(counts)]).astype(np.int64), col, val)


def doRandomWalks(similarities: SparseRows, rws: RandomWalkSettings, rank: int = 0, world: int = 1, group=None, walks_fn=None) -> SparseRows:
    """utils::doRandomWalks; with world > 1 the start points are sharded over the ranks (the transition matrix is
    replicated, every rank walks its slab of start points, one gather of the ragged result)."""
    from . import capi

    if world > 1:
        q0, q1 = shard_rows(similarities.n, world)[rank]
        fn = walks_fn or (lambda m, q0, q1: capi.random_walks_slab(m.rowptr, m.col, m.val, rws.numRandomWalks, rws.singleWalkLength, q0, q1,
                                                                    int(rws.importanceWeighting), rws.pruneValue, rws.pruneSteps, rws.randomSeed,
                                                                    rws.removeDiagonal, rws.normalize))
        rp, c, v = fn(similarities, q0, q1)
        return gather_csr(SparseRows(rp, c, v), similarities.n, group)
    rp, c, v = capi.random_walks(similarities.rowptr, similarities.col, similarities.val, rws.numRandomWalks, rws.singleWalkLength,
                                 int(rws.importanceWeighting), rws.pruneValue, rws.pruneSteps, rws.randomSeed, rws.removeDiagonal, rws.normalize)
    return SparseRows(rp, c, v)


def createSimilarities(matrix: SparseRows, k: int, prune_val: float = 1e-4, variant: int = 0, comp_sizes=None, rank: int = 0, world: int = 1, group=None,
                       sims_fn=None):
    """utils::createSimilarities (sph/utils/SparseMatrixAlgorithms.cpp:963-1378): (cols [n, k] padded 0xFFFFFFFF, vals [n, k], nnz [n]).
    With world > 1 the OUTPUT ROWS are sharded over the ranks (the random-walk matrix is replicated, every rank computes
    the rows of its slab against the whole matrix, one gather of the fixed-stride result)."""
    from . import capi

    if world > 1:
        r0, r1 = shard_rows(matrix.n, world)[rank]
        fn = sims_fn or (lambda m, r0, r1: capi.create_similarities_slab(m.rowptr, m.col, m.val, k, r0, r1, prune_val, variant, comp_sizes))
        cols, vals, nnz = fn(matrix, r0, r1)
        cols = gather_rows(np.ascontiguousarray(cols, np.uint32).view(np.int32), matrix.n, group).view(np.uint32)
        vals = gather_rows(np.ascontiguousarray(vals, np.float32), matrix.n, group)
        nnz = gather_rows(np.ascontiguousarray(nnz, np.int32), matrix.n, group)
        return cols, vals, nnz
    return capi.create_similarities(matrix.rowptr, matrix.col, matrix.val, k, prune_val, variant, comp_sizes)


def mergeNodesRandomWalks(walks: SparseRows, num_merged: int, parents, norm: bool = True, weight_by_size: bool = True) -> SparseRows:
    from . import capi

    rp, c, v = capi.rw_merge(walks.rowptr, walks.col, walks.val, parents, num_merged, norm, weight_by_size)
    return SparseRows(rp, c, v)


def mergeGraphNodes(graph, num_merged: int, parents) -> RaggedGraph:
    """utils::mergeGraphNodes (sph/utils/SparseMatrixAlgorithms.cpp:486-561)."""
    from . import capi

    rowptr = getattr(graph, "rowptr", None)
    rp, ids, dist = capi.merge_graph_nodes(graph.knnIndices, graph.knnDistances, parents, num_merged, rowptr=rowptr)
    return RaggedGraph(rp, ids, dist)


def mergeNodesDataDistances(matrix: SparseRows, num_merged: int, parents) -> SparseRows:
    """utils::mergeNodesDataDistances (sph/utils/SparseMatrixAlgorithms.cpp:443-484)."""
    from . import capi

    rp, c, v = capi.merge_data_distances(matrix.rowptr, matrix.col, matrix.val, parents, num_merged)
    return SparseRows(rp, c, v)


def symmetrizeGraph(graph) -> RaggedGraph:
    """utils::symmetrizeGraph (sph/utils/GraphUtils.cpp:197-317)."""
    from . import capi

    rp, ids, dist = capi.symmetrize_graph(graph.knnIndices, graph.knnDistances, rowptr=getattr(graph, "rowptr", None))
    return RaggedGraph(rp, ids, dist, symmetric=True)


def euclidDistance(data: np.ndarray, comp_ptr, comp_pix, pairs_a, pairs_b) -> np.ndarray:
    """Batched utils::euclidDistance (sph/utils/Similarities.cpp:414-483) over component pairs."""
    from . import capi

    return capi.euclid_hausdorff(data, comp_ptr, comp_pix, pairs_a, pairs_b)


# ---- the reference's stage-cache blobs (sph/utils/FileIO.{hpp,cpp}): LZ4 files this back-end writes and reads so
# that its results can be dropped into the reference's cache directory (and the reference's loaded here)
def writeCompressedGraphToBinary(file_name_base: str, graph) -> None:
    """utils::writeCompressedGraphToBinary (sph/utils/FileIO.cpp:586-618)."""
    from . import capi

    capi.cache_write_graph(file_name_base, graph.knnIndices, graph.knnDistances, rowptr=getattr(graph, "rowptr", None), symmetric=bool(graph.symmetric))


def loadCompressedGraphFromBinary(file_name_base: str) -> RaggedGraph:
    """utils::loadCompressedGraphFromBinary (sph/utils/FileIO.cpp:895-937)."""
    from . import capi

    rp, ids, dist, sym = capi.cache_read_graph(file_name_base)
    return RaggedGraph(rp, ids, dist, symmetric=sym)


def writeCompressedSparseMatHDIToBinary(path: str, matrix: SparseRows) -> None:
    """utils::writeCompressedSparseMatHDIToBinary (sph/utils/FileIO.cpp:620-799): a P matrix, e.g. `_lls_probDist<level>.cache`."""
    from . import capi

    capi.cache_write_prob(path, matrix.rowptr, matrix.col, matrix.val)


def loadCompressedSparseMatHDIFromBinary(path: str) -> SparseRows:
    """utils::loadCompressedSparseMatHDIFromBinary (sph/utils/FileIO.cpp:1274-1460)."""
    from . import capi

    return SparseRows(*capi.cache_read_prob(path))


def writeCompressedVecToBinary(path: str, vec: np.ndarray) -> None:
    """utils::writeCompressedVecToBinary<T> (sph/utils/FileIO.hpp:279-291), e.g. the connected-component labels."""
    from . import capi

    capi.cache_write_vec(path, vec)


def loadCompressedVecFromBinary(path: str, dtype) -> np.ndarray:
    """utils::loadCompressedVecFromBinary<T> (sph/utils/FileIO.hpp:523-640)."""
    from . import capi

    return capi.cache_read_vec(path, dtype)


def geodesicDistance(graph, comp_ptr, comp_pix, pairs_a, pairs_b, labels=None) -> np.ndarray:
    """Batched utils::geodesicDistance (sph/utils/Similarities.cpp:234-309) over component pairs: one single-source search on the
    device per point of every first component (sphb_geodesic_hausdorff)."""
    from . import capi

    return capi.geodesic_hausdorff(graph.knnIndices, graph.knnDistances, comp_ptr, comp_pix, pairs_a, pairs_b, rowptr=getattr(graph, "rowptr", None),
                                   labels=labels)


def subsampleComponentPairs(comp_ptr, comp_pix, pairs_a, pairs_b, num_samples: int | None, rng=None):
    """The sub-sampling branch of geodesicDistance / euclidDistance (sph/utils/Similarities.cpp:283-303, :455-474): a pair with more
    than numGeodesicSamples^2 point combinations is evaluated on numGeodesicSamples points drawn WITH replacement from either
    component (the reference seeds from random_device, so its draws are not reproducible; `rng` makes ours so).  Returns
    (comp_ptr, comp_pix, pairs_a, pairs_b) with the sampled lists appended as extra components."""
    comp_ptr = np.asarray(comp_ptr, np.int64)
    comp_pix = np.asarray(comp_pix, np.int64)
    pairs_a = np.array(pairs_a, np.int64)
    pairs_b = np.array(pairs_b, np.int64)
    if num_samples is None:
        return comp_ptr, comp_pix, pairs_a, pairs_b
    sizes = np.diff(comp_ptr)
    big = np.nonzero(sizes[pairs_a] * sizes[pairs_b] > num_samples * num_samples)[0]
    if len(big) == 0:
        return comp_ptr, comp_pix, pairs_a, pairs_b
    rng = rng or np.random.default_rng()
    extra = []
    ncomp = len(comp_ptr) - 1
    for p in big:
        for side in (pairs_a, pairs_b):
            c = side[p]
            lst = comp_pix[comp_ptr[c]:comp_ptr[c + 1]]
            extra.append(lst[rng.integers(0, len(lst), num_samples)])
            side[p] = ncomp
            ncomp += 1
    comp_pix = np.concatenate([comp_pix] + extra)
    comp_ptr = np.concatenate([comp_ptr, comp_ptr[-1] + num_samples * np.arange(1, len(extra) + 1, dtype=np.int64)])
    return comp_ptr, comp_pix, pairs_a, pairs_b


# ---- ImageHierarchy (sph/ImageHierarchy.hpp:46, .cpp:113-608) — the HOST merge loop ---------------
class NeighConnection(enum.IntEnum):
    FOUR = 0
    EIGHT = 1


class RandomWalkHandling(enum.IntEnum):
    MERGE_RW_ONLY = 0
    MERGE_RW_NEW_WALKS = 1
    MERGE_RW_NEW_WALKS_AND_KNN = 2
    MERGE_DATA_NEW_WALKS = 3


class RandomWalkReduction(enum.IntEnum):
    NONE = 0
    PROPORTIONAL_COMPONENT_REDUCTION = 1
    PROPORTIONAL_HALF = 2
    PROPORTIONAL_DOUBLE = 3
    CONSTANT = 4
    CONSTANT_LOW = 5
    CONSTANT_HIGH = 6


class NormType(enum.IntEnum):
    ONEDIM = 0
    TWODIM = 1


FLT_MAX = float(np.finfo(np.float32).max)


@dataclass
class ImageHierarchySettings:  # sph/utils/Settings.hpp:145-166
    componentSim: ComponentSim = ComponentSim.NEIGH_OVERLAP
    neighborConnection: NeighConnection = NeighConnection.FOUR
    mergeMultiple: bool = False
    usePercentile: bool = True
    maxDist: float = 0.0
    minNumComp: int = 1
    minReduction: float = 99.99
    numGeodesicSamples: int | None = None
    maxLevels: int = -1
    rwHandling: RandomWalkHandling = RandomWalkHandling.MERGE_RW_ONLY
    rwReduction: RandomWalkReduction = RandomWalkReduction.PROPORTIONAL_COMPONENT_REDUCTION
    normKnnDistances: NormalizationScheme = NormalizationScheme.TSNE
    rwWeightMergeBySize: bool = True
    rwNormSim: NormType = NormType.ONEDIM
    rwRemoveSelfSimAfterMerging: bool = True
    componentLabels: np.ndarray | None = None  # connected-component label per point (GEO_*: sph/ComputeHierarchy.cpp:143)

    def isAlwaysMerge(self) -> bool:
        return self.maxDist == -1.0


class DeviceOps:
    """The compute calls the merge loop makes, bound to the C ABI (the product path).  Tests substitute an object with the
    same methods backed by the CPU oracle to pin the loop against the reference's hierarchy goldens without a GPU."""

    def prow(self, graph: Graph, scheme: NormalizationScheme) -> SparseRows:
        return normalizeKnnDistances(graph, scheme)

    def random_walks(self, sims: SparseRows, rws: RandomWalkSettings) -> SparseRows:
        return doRandomWalks(sims, rws)

    def rw_merge(self, walks: SparseRows, parents, n_merged: int, norm: bool, weight_by_size: bool) -> SparseRows:
        return mergeNodesRandomWalks(walks, n_merged, parents, norm, weight_by_size)

    def rw_post_merge(self, merged: SparseRows, remove_diagonal: bool, norm) -> SparseRows:
        from . import capi

        return SparseRows(*capi.rw_post_merge(merged.rowptr, merged.col, merged.val, remove_diagonal, 1 if norm == NormType.ONEDIM else 2))

    def pair_sims(self, walks: SparseRows, pa, pb, mode: int) -> np.ndarray:
        from . import capi

        return capi.pair_sims(walks.rowptr, walks.col, walks.val, pa, pb, mode)

    def merge_graph_nodes(self, graph: RaggedGraph, parents, n_merged: int) -> RaggedGraph:
        return mergeGraphNodes(graph, n_merged, parents)

    def prow_ragged(self, graph: RaggedGraph, scheme: NormalizationScheme) -> SparseRows:
        from . import capi

        return SparseRows(*capi.prow_graph_ragged(graph.rowptr, graph.knnIndices, graph.knnDistances, 0.0, scheme == NormalizationScheme.LINEAR))

    def neighbor_overlap(self, knn_ids, comp_ptr, comp_pix, pa, pb) -> np.ndarray:
        from . import capi

        return capi.neighbor_overlap(knn_ids, comp_ptr, comp_pix, pa, pb)

    def euclid_hausdorff(self, data, comp_ptr, comp_pix, pa, pb) -> np.ndarray:
        from . import capi

        return capi.euclid_hausdorff(data, comp_ptr, comp_pix, pa, pb)

    def geodesic_hausdorff(self, graph, comp_ptr, comp_pix, pa, pb, labels=None) -> np.ndarray:
        # the data graph goes to the device once per hierarchy (what _hierarchy.bgraph is to the reference, ImageHierarchy.cpp:427-431)
        from . import capi

        key = (id(graph), id(labels))
        if getattr(self, "_geo_key", None) != key:
            if getattr(self, "_geo_graph", None) is not None:
                self._geo_graph.close()
            self._geo_graph = capi.GeodesicGraph(graph.knnIndices, graph.knnDistances, rowptr=getattr(graph, "rowptr", None), labels=labels)
            self._geo_key = key
        return self._geo_graph.hausdorff(comp_ptr, comp_pix, pa, pb)


def pixelNeighborIDs(num_cols: int, num_rows: int, conn: NeighConnection, pid: int) -> list[int]:
    """sph/utils/ImageHelper.cpp:8-29 with the direction tables of ImageHelper.hpp:13-19 (the order matters at level 0:
    ties in distance keep it)."""
    if conn == NeighConnection.EIGHT:
        rd, cd = (-1, -1, -1, 0, 0, 1, 1, 1), (-1, 0, 1, -1, 1, -1, 0, 1)
    else:
        rd, cd = (-1, 0, 1, 0), (0, 1, 0, -1)
    r, c = divmod(pid, num_cols)
    out = []
    for dr, dc in zip(rd, cd):
        nr, nc = r + dr, c + dc
        if 0 <= nr < num_rows and 0 <= nc < num_cols:
            out.append(nr * num_cols + nc)
    return out


def labelWeakComponents(n: int, edges) -> tuple[int, np.ndarray]:
    """utils::labelGraphWeakComponents (sph/utils/GraphUtils.cpp:371-392): boost::strong_components on the symmetrised
    graph numbers the components in the order their lowest vertex is reached."""
    parent = list(range(n))

    def find(a):
        while parent[a] != a:
            parent[a] = parent[parent[a]]
            a = parent[a]
        return a

    for a, b in edges:
        ra, rb = find(a), find(b)
        if ra != rb:
            parent[max(ra, rb)] = min(ra, rb)
    labels = np.empty(n, np.int64)
    seen: dict[int, int] = {}
    for v in range(n):
        r = find(v)
        if r not in seen:
            seen[r] = len(seen)
        labels[v] = seen[r]
    return len(seen), labels


def computeQuantile(data: np.ndarray, quantile: float, ignore=(), interpolation: int = 0) -> float:
    """sph/utils/Math.cpp:133-165."""
    d = np.asarray(data, np.float32)
    for iv in ignore:
        d = d[d != np.float32(iv)]
    d = np.sort(d)
    rank = np.float32(quantile) * np.float32(len(d) - 1)
    lo, hi = int(np.floor(rank)), int(np.ceil(rank))
    if lo == hi:
        return float(d[lo])
    frac = np.float32(rank - np.float32(lo))
    if interpolation == 1:
        return float(d[lo] + (d[hi] - d[lo]) * frac)
    return float(np.float32(0.5) * (d[lo] + d[hi]))


def _seq_sum_f32(v: np.ndarray) -> np.float32:  # Eigen's SparseVector::sum(): fp32, storage order
    return np.cumsum(v, dtype=np.float32)[-1] if len(v) else np.float32(0)


def removeDiagonalAndNormalize(m: SparseRows, remove_diagonal: bool, norm: NormType) -> SparseRows:
    """Hierarchy::updateRandomWalks after the merge (sph/utils/Hierarchy.cpp:287-307): removeDiagonalElements(keepSingleEntry,
    SparseMatrixAlgorithms.cpp:704-718), then per-row (ONEDIM) or whole-matrix (TWODIM) L1 normalisation."""
    rowptr, cols, vals = [0], [], []
    for i in range(m.n):
        c, v = m.row(i)
        if remove_diagonal and len(c) > 1:
            keep = (c != i) & (v != 0)  # removeElement: coeffRef(i) = 0, then prune(0, eps) drops every exact zero of the row
            c, v = c[keep], v[keep]
        cols.append(c)
        vals.append(np.array(v, np.float32))
        rowptr.append(rowptr[-1] + len(c))
    if norm == NormType.ONEDIM:
        vals = [v / _seq_sum_f32(v) if len(v) else v for v in vals]
    else:
        tot = np.float32(0)
        for v in vals:
            tot = np.float32(tot + _seq_sum_f32(v))
        vals = [v / tot for v in vals]
    return SparseRows(np.array(rowptr, np.int64), np.concatenate(cols).astype(np.uint32) if cols else np.zeros(0, np.uint32),
                      np.concatenate(vals).astype(np.float32) if vals else np.zeros(0, np.float32))


class ImageHierarchy:
    """Borůvka-style superpixel merge loop (sph/ImageHierarchy.cpp:113-591) with the Hierarchy bookkeeping of
    sph/utils/Hierarchy.cpp:134-390.  It stays on the host (BASELINE.json north_star); every data-parallel step it takes —
    level-0 P rows, random walks, walk merges, component distances of all (component, spatial neighbour) pairs of a level —
    is one batched call into `ops` (DeviceOps = the CUDA path through the C ABI)."""

    def __init__(self, knn_graph: Graph, data: np.ndarray | None, num_rows: int, num_cols: int, ops=None):
        self._graph = knn_graph
        self._data = data
        self._rows, self._cols = int(num_rows), int(num_cols)
        self._ops = ops or DeviceOps()
        self.ihs = ImageHierarchySettings()
        self.rws = RandomWalkSettings()
        self._clear()

    def _clear(self):
        self.numComponents: list[int] = []
        self.parents: list[np.ndarray] = []
        self.children: list[list[list[int]]] = []
        self.spatialNeighbors: list[list[list[int]]] = []
        self.pixelComponents: list[np.ndarray] = []
        self.randomWalks: list[SparseRows] = []
        self.mergedDataGraphs: list[RaggedGraph] = []
        self.notMergedNodes: list[list[int]] = []
        self.reductionRates: list[float] = []
        self.rwRandomWalkLengths: list[int] = []
        self.dataLevelProbdist: SparseRows | None = None

    # -- accessors named after the reference's
    def getNumLevels(self) -> int:
        return len(self.numComponents)

    def pixelComponentsOn(self, level: int) -> np.ndarray:
        return self.pixelComponents[level]

    def parentsOn(self, level: int) -> np.ndarray:
        return self.parents[level]

    def childrenOn(self, level: int):
        return self.children[level - 1]

    def spatialNeighborsOn(self, level: int):
        return self.spatialNeighbors[level - 1]

    def getDataLevelProbdist(self) -> SparseRows:
        return self.dataLevelProbdist

    def getSpatialNeighbors(self, level: int, cid: int) -> list[int]:  # ImageHierarchy.cpp:594-605
        if level == 0:
            return pixelNeighborIDs(self._cols, self._rows, self.ihs.neighborConnection, cid)
        return self.spatialNeighbors[level - 1][cid]

    def representedDataPoints(self, level: int):
        """(comp_ptr, comp_pix) of every component of `level` (Hierarchy::getRepresentedDataPoints, Hierarchy.cpp:27-73)."""
        n = self.numComponents[level]
        if level == 0:
            return np.arange(n + 1, dtype=np.int64), np.arange(n, dtype=np.int64)
        lists = [list(ch) for ch in self.children[level - 1]]
        for lower in range(level - 1, 0, -1):
            lists = [[g for c in comp for g in self.children[lower - 1][c]] for comp in lists]
        ptr = np.concatenate([[0], np.cumsum([len(c) for c in lists])]).astype(np.int64)
        return ptr, np.array([p for c in lists for p in c], np.int64)

    def _uses_walks(self) -> bool:
        return self.ihs.componentSim in (ComponentSim.NEIGH_WALKS, ComponentSim.NEIGH_WALKS_SINGLE_OVERLAP, ComponentSim.GEO_WALKS)

    # -- compute
    def compute(self, ihs: ImageHierarchySettings | None = None, rws: RandomWalkSettings | None = None) -> None:
        if ihs is not None:
            self.ihs = ihs
        if rws is not None:
            self.rws = rws
        self._clear()
        if self.ihs.normKnnDistances not in (NormalizationScheme.TSNE, NormalizationScheme.LINEAR) and self._uses_walks():
            log.error("ImageHierarchy::compute: %s rows are not built (UMAP rows are a 'next' item)", self.ihs.normKnnDistances.name)
            return
        try:
            self._computePreparations()
            self._computeBoruvkaHierarchy()
        except Exception as e:  # reference convention: log + return
            log.error("ImageHierarchy::compute failed: %s", e)
            self._clear()

    def _computePreparations(self):  # ImageHierarchy.cpp:149-190
        if self.ihs.normKnnDistances in (NormalizationScheme.TSNE, NormalizationScheme.LINEAR):
            if isinstance(self._graph, RaggedGraph):  # the connected / symmetrised graph of the GEO_* configurations
                self.dataLevelProbdist = self._ops.prow_ragged(self._graph, self.ihs.normKnnDistances)
            else:
                self.dataLevelProbdist = self._ops.prow(self._graph, self.ihs.normKnnDistances)
        n = self._graph.numPoints
        self.numComponents = [n]
        self.pixelComponents = [np.arange(n, dtype=np.int64)]
        if self._uses_walks():
            self.rwRandomWalkLengths.append(self.rws.singleWalkLength)
            self.randomWalks.append(self._ops.random_walks(self.dataLevelProbdist, self.rws))
            if self.ihs.rwHandling == RandomWalkHandling.MERGE_DATA_NEW_WALKS:  # keep the data distances for later merging (.cpp:176-177)
                g = self._graph
                if isinstance(g, RaggedGraph):
                    self.mergedDataGraphs.append(g)
                else:
                    k = g.knnIndices.shape[1]
                    self.mergedDataGraphs.append(RaggedGraph(np.arange(n + 1, dtype=np.int64) * k, g.knnIndices.reshape(-1), g.knnDistances.reshape(-1)))

    def computeDistances(self, level: int):
        """ImageHierarchy::computeDistances (.cpp:192-249): rows [self, spatial neighbours...], componentDistance for every
        pair in ONE batched call, rows stable-sorted by distance."""
        n = self.numComponents[level]
        neigh = [[c] + list(self.getSpatialNeighbors(level, c)) for c in range(n)]
        rowptr = np.concatenate([[0], np.cumsum([len(r) for r in neigh])]).astype(np.int64)
        ids = np.array([v for r in neigh for v in r], np.int64)
        pa = np.repeat(np.arange(n, dtype=np.int64), np.diff(rowptr))
        sim = self.ihs.componentSim
        if sim == ComponentSim.NEIGH_OVERLAP:
            ptr, pix = self.representedDataPoints(level)
            dist = np.float32(1) - self._ops.neighbor_overlap(self._graph.knnIndices, ptr, pix, pa, ids)
        elif sim in (ComponentSim.NEIGH_WALKS, ComponentSim.NEIGH_WALKS_SINGLE_OVERLAP):
            mode = 1 if sim == ComponentSim.NEIGH_WALKS_SINGLE_OVERLAP else 0
            dist = np.float32(1) - self._ops.pair_sims(self.randomWalks[level], pa, ids, mode)
        elif sim == ComponentSim.EUCLID_CENTROID:
            ptr, pix = self.representedDataPoints(level)
            sptr, spix, spa, spb = subsampleComponentPairs(ptr, pix, pa, ids, self.ihs.numGeodesicSamples)
            dist = self._ops.euclid_hausdorff(self._data, sptr, spix, spa, spb)
        elif sim in (ComponentSim.GEO_CENTROID, ComponentSim.GEO_WALKS):  # geodesicDistance (Similarities.cpp:142,146)
            ptr, pix = self.representedDataPoints(level)
            sptr, spix, spa, spb = subsampleComponentPairs(ptr, pix, pa, ids, self.ihs.numGeodesicSamples)
            dist = self._ops.geodesic_hausdorff(self._graph, sptr, spix, spa, spb, self.ihs.componentLabels)
        else:
            raise NotImplementedError(sim)
        dist = np.asarray(dist, np.float32).copy()
        dist[pa == ids] = 0  # componentDistance: same id -> 0 (Similarities.cpp:127-128)
        for c in range(n):  # synchronizedSort, stable (Algorithms.hpp:57-93)
            b, e = rowptr[c], rowptr[c + 1]
            order = np.argsort(dist[b:e], kind="stable")
            dist[b:e] = dist[b:e][order]
            ids[b:e] = ids[b:e][order]
        return RaggedGraph(rowptr, ids, dist)

    def _connectMostSimilarComponents(self, level: int):
        """connectMostSimilarComponents + mergeMinBelow / mergeAllBelow (.cpp:260-407): the merge edges of one level."""
        d = self.computeDistances(level)
        s = self.ihs
        thresh = np.float32(FLT_MAX)
        if s.maxDist > 0:
            thresh = np.float32(s.maxDist)
            if s.usePercentile:
                q = computeQuantile(d.knnDistances, s.maxDist, (0.0, -1.0, FLT_MAX))
                thresh = np.float32(q if q >= 0 else FLT_MAX)
        edges = []
        rng = np.random.default_rng()
        for c in range(d.numPoints):
            ids, dist = d.getNeighbors(c), d.getDistances(c)
            if s.mergeMultiple:
                merge = []
                for nid, dv in zip(ids[1:], dist[1:]):
                    if dv < thresh:
                        merge.append(int(nid))
                    else:
                        break
                if not merge:
                    merge = [int(ids[rng.integers(1, len(ids))])] if s.isAlwaysMerge() and len(ids) > 1 else [c]
                edges.extend((c, m) for m in merge)
            else:
                merge, dmax = c, thresh
                if len(ids) > 1 and dist[1] < dmax:  # distances are sorted: only the first neighbour can be the minimum
                    dmax, merge = dist[1], int(ids[1])
                if dmax == np.float32(FLT_MAX) and s.isAlwaysMerge() and len(ids) > 1:
                    merge = int(ids[rng.integers(1, len(ids))])
                if merge != c:
                    edges.append((c, merge))
        return edges

    def _next_walk_length(self) -> None:  # .cpp:489-533
        s = self.ihs
        cur = self.rwRandomWalkLengths[-1]
        rate = np.float32(self.reductionRates[-1]) / np.float32(100)
        r = {RandomWalkReduction.NONE: np.float32(1), RandomWalkReduction.PROPORTIONAL_COMPONENT_REDUCTION: rate,
             RandomWalkReduction.PROPORTIONAL_DOUBLE: rate * np.float32(2), RandomWalkReduction.PROPORTIONAL_HALF: rate * np.float32(0.5),
             RandomWalkReduction.CONSTANT: np.float32(0.5), RandomWalkReduction.CONSTANT_LOW: np.float32(0.75),
             RandomWalkReduction.CONSTANT_HIGH: np.float32(0.25)}[s.rwReduction]
        r = min(max(r, np.float32(0)), np.float32(1))
        nxt = int(np.float32(r * np.float32(cur)))
        lo, hi = self.rws.minimumSingleWalkLength, self.rwRandomWalkLengths[0]
        nxt = lo if nxt < lo else (hi if hi < nxt else nxt)  # std::clamp as libstdc++ evaluates it (lo may exceed hi)
        self.rwRandomWalkLengths.append(nxt)

    def _computeBoruvkaHierarchy(self):  # .cpp:409-591
        s = self.ihs
        num_trees = self.numComponents[0]
        level = 0
        while num_trees > s.minNumComp:
            if s.maxLevels >= 0 and level >= s.maxLevels:
                break
            n_cur = self.numComponents[level]
            edges = self._connectMostSimilarComponents(level)
            n_next, labels = labelWeakComponents(n_cur, edges)
            self.reductionRates.append(float(np.float32(100) * np.float32(n_next) / np.float32(n_cur)))
            rr = self.reductionRates
            if rr[-1] == 100.0 or (len(rr) > 2 and rr[-1] > s.minReduction and rr[-2] > s.minReduction):
                break
            if self._uses_walks() and s.rwHandling != RandomWalkHandling.MERGE_RW_ONLY:
                self._next_walk_length()
            self._addLevel(n_next, labels)
            num_trees = n_next
            level += 1

    def _addLevel(self, n_next: int, labels: np.ndarray):  # Hierarchy::addLevel (Hierarchy.cpp:134-248)
        n_pix = self.numComponents[0]
        cur_level = len(self.numComponents) - 1
        self.numComponents.append(int(n_next))
        self.parents.append(labels.astype(np.uint64))
        self.pixelComponents.append(labels[self.pixelComponents[cur_level]].astype(np.int64))
        children = [[] for _ in range(n_next)]
        for c, lab in enumerate(labels):
            children[int(lab)].append(c)
        self.children.append(children)
        self.notMergedNodes.append([i for i, ch in enumerate(children) if len(ch) == 1])
        pix = self.pixelComponents[-1]
        sp = [set() for _ in range(n_next)]
        for p in range(n_pix):
            for q in pixelNeighborIDs(self._cols, self._rows, self.ihs.neighborConnection, p):
                if pix[q] != pix[p]:
                    sp[int(pix[p])].add(int(pix[q]))
        self.spatialNeighbors.append([sorted(x) for x in sp])
        if self._uses_walks():
            self._updateRandomWalks(n_next)

    def _updateRandomWalks(self, n_next: int):  # Hierarchy::updateRandomWalks (Hierarchy.cpp:250-390)
        s = self.ihs
        if s.rwHandling == RandomWalkHandling.MERGE_DATA_NEW_WALKS:  # Hierarchy.cpp:309-339: merge the data graph, rows from its distances
            g = self._ops.merge_graph_nodes(self.mergedDataGraphs[-1], self.parents[-1], n_next)
            merged = self._ops.prow_ragged(g, s.normKnnDistances)
            self.mergedDataGraphs.append(g)
        else:
            merged = self._ops.rw_merge(self.randomWalks[-1], self.parents[-1], n_next, False, s.rwWeightMergeBySize)
            remove = s.rwRemoveSelfSimAfterMerging and merged.n > 1 and s.rwHandling != RandomWalkHandling.MERGE_RW_ONLY
            merged = self._ops.rw_post_merge(merged, remove, s.rwNormSim)
        if s.rwHandling == RandomWalkHandling.MERGE_RW_ONLY:
            out = merged
        else:
            rws = RandomWalkSettings(**{**self.rws.__dict__, "singleWalkLength": self.rwRandomWalkLengths[-1]})
            out = self._ops.random_walks(merged, rws)
        if out.n == 1 and len(out.col) == 0:  # a single top node keeps its self-similarity
            out = SparseRows(np.array([0, 1], np.int64), np.zeros(1, np.uint32), np.ones(1, np.float32))
        self.randomWalks.append(out)
