/*
 * sphb.h — C ABI of the B200-native back-end for the SPH hot path.
 *
 * The reference (alxvth/SPH @ 8d54e737) has no FFI: its hot path is a set of
 * C++ free functions in namespace sph::utils that call faiss / HDILib / Eigen
 * on host containers.  This header is the seam a maintainer binds instead
 * (INTEGRATION.md shows the shim for every entry point).  Every function cites
 * the reference interface it replaces.
 *
 * Conventions
 *   - plain pointers and sizes only; the caller owns every buffer;
 *   - entry points without a suffix take HOST pointers (pageable or pinned) and
 *     do their own H2D / D2H copies; `_dev` variants take DEVICE pointers and a
 *     cudaStream_t (passed as void*) and never touch the host — they are what a
 *     pipeline that keeps data resident in HBM calls;
 *   - return 0 on success, non-zero on error; sphb_last_error() gives the text.
 *     The reference's convention is "log + early return, outputs untouched"
 *     (sph/LevelSimilarities.cpp:123-142); the shim maps non-zero to that;
 *   - there is NO CPU fallback: without a CUDA device every call fails with
 *     SPHB_ERR_NO_DEVICE;
 *   - single caller at a time per process (the reference facade is sequential,
 *     sph/ComputeHierarchy.cpp:262-276).
 *   - threading: a context (device, stream, last error, last-call statistics) belongs to the HOST THREAD that created it;
 *     different threads may call concurrently (each on its own context), one thread's calls are sequential.  Transient device
 *     buffers are cached per stream between calls (also for a caller's stream passed to a _dev entry point): call
 *     sphb_shutdown() in the thread that used a stream before destroying that stream.
 *   - size limits (exceeding one returns SPHB_ERR_UNSUPPORTED / SPHB_ERR_CAPACITY with a message, never a truncated result):
 *       kNN: n < 2^31 points on the tensor-core path (2^32 - 1 on the SIMT path), k <= 864 on the tensor-core path
 *            (next_pow2(k + 160) <= 1024), any d >= 1 (d < 8 or n < 1024 run the exact SIMT kernel);
 *       P rows: k <= 1024 per row (graph, ragged and in-place variants);
 *       random walks: numRandomWalks * singleWalkLength <= 8192 events per start point, FIRST_VISIT walk length <= 64;
 *       createSimilarities / level kNN: k <= 992, n < 2^32; findTopK / findBottomK: rows of at most 4096 entries;
 *       every sort-based pass (symmetrize, rw_merge, symmetrize_graph, merge_*): at most 2^31 - 1 records per call
 *       (symmetrize expands to 2 * nnz records);
 *   - sparse matrices cross the boundary as CSR: int64 rowptr[n+1],
 *     uint32 col[nnz] (column-sorted within a row), float val[nnz] — the flat
 *     form of SparseMatSPH / SparseMatHDI (sph/utils/CommonDefinitions.hpp:87-93).
 */
#ifndef SPHB_H
#define SPHB_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define SPHB_API __attribute__((visibility("default")))
#else
#define SPHB_API
#endif

enum {
    SPHB_OK = 0,
    SPHB_ERR_NO_DEVICE = 1,
    SPHB_ERR_INVALID = 2,
    SPHB_ERR_CUDA = 3,
    SPHB_ERR_CAPACITY = 4,
    SPHB_ERR_UNSUPPORTED = 5
};

/* faiss::MetricType values used at the reference call sites
 * (sph/NearestNeighbors.cpp:81-96): METRIC_INNER_PRODUCT = 0, METRIC_L2 = 1. */
enum { SPHB_METRIC_INNER_PRODUCT = 0, SPHB_METRIC_L2 = 1,
       /* KnnMetric::COSINE (sph/utils/Settings.hpp; NearestNeighbors::setMetric + initFaissData, sph/NearestNeighbors.cpp:81-96, 208-222):
        * not a faiss value — the back-end L2-normalises a device copy of the rows and searches it with the inner product */
       SPHB_METRIC_COSINE = 2 };

/* sph::utils::ImportanceWeighting (sph/utils/Settings.hpp, used at
 * sph/utils/SparseMatrixAlgorithms.cpp:137-161). */
enum { SPHB_W_CONSTANT = 0, SPHB_W_LINEAR = 1, SPHB_W_NORMAL = 2, SPHB_W_ONLYLAST = 3, SPHB_W_FIRST_VISIT = 4 };

/* ---- lifetime ---------------------------------------------------------- */
SPHB_API int sphb_init(int device);          /* selects the device, creates the stream; idempotent */
SPHB_API void sphb_shutdown(void);           /* releases device scratch */
SPHB_API const char* sphb_last_error(void);
SPHB_API int sphb_version(void);
/* number of kernels this library has launched since sphb_init / the last reset */
SPHB_API int64_t sphb_launch_count(void);
SPHB_API void sphb_reset_launch_count(void);
/* the stream all host-pointer entry points run on (cudaStream_t) */
SPHB_API void* sphb_stream(void);

/* ---- Stage A: exact kNN graph ------------------------------------------ */
/*
 * Replaces utils::computeButeForce / utils::computeIndexFlat
 *   (sph/utils/Knn.hpp:10-11, sph/utils/Knn.cpp:56-136)
 * fused with the post-passes NearestNeighbors::compute applies right after
 *   (sph/NearestNeighbors.cpp:152-167: ensureFloatingPointIntegrity :233-242,
 *    squareRootOfAllDistances :224-231, ensureClosestPointIsSelf
 *    sph/utils/GraphUtils.cpp:23-95).
 * x: n*d float32 point-major.  ids/dist: n*k, caller-owned (the reference
 * pre-fills -1 / -1.0f, sph/NearestNeighbors.cpp:113-118; slots that cannot be
 * filled because n < k keep that value).  Row p is ordered by (distance, index),
 * distances are the fp64-accumulated squared L2 rounded once to fp32, then
 * `<= FLT_EPSILON -> 0`, then sqrt unless l2_squared, then self moved first.
 */
SPHB_API int sphb_knn_exact(const float* x, int64_t n, int d, int k, int metric, int l2_squared,
                            int64_t* ids, float* dist);
/* Same for query rows [q0, q1) against the full database (multi-GPU shard):
 * ids/dist are (q1-q0)*k. */
SPHB_API int sphb_knn_exact_slab(const float* x, int64_t n, int d, int k, int metric, int l2_squared,
                                 int64_t q0, int64_t q1, int64_t* ids, float* dist);
SPHB_API int sphb_knn_exact_dev(const float* d_x, int64_t n, int d, int k, int metric, int l2_squared,
                                int64_t q0, int64_t q1, int64_t* d_ids, float* d_dist, void* stream);
/* faiss::fvec_renorm_L2 on a copy, the COSINE preparation of NearestNeighbors::initFaissData (sph/NearestNeighbors.cpp:208-222):
 * out[i] = x[i] / sqrt(fp32 sum of x[i][c]^2 in column order); all-zero rows are copied.  (sphb_knn_exact* with
 * SPHB_METRIC_COSINE does this on the device by itself; this entry point serves callers that keep the normalised copy.) */
SPHB_API int sphb_renorm_l2(const float* x, int64_t n, int d, float* out);
/* kNN statistics of the last call: [0] path (0 SIMT, 1 tcgen05 streaming, 2 tcgen05 resident MH=1,
 * 3 tcgen05 resident MH=2, 4 tcgen05 CTA pair (cta_group::2)), [1] candidate-list compactions, [2] candidates accepted by the filter,
 * [3] executed GEMM K extent per pair of the main sweep, [4] rows recomputed by the exact SIMT kernel
 * (candidate band overflow), [5] query tiles, [6] fp16 products per pair of the main sweep (1 or 3),
 * [7] rows re-swept with the 3-product filter after overflowing the 1-product band. */
SPHB_API int sphb_knn_last_stats(int64_t* out8);
/* device time of the last tensor-core kNN call in ms (CUDA events on its stream): [0] operand
 * preparation, [1] the tcgen05 sweep kernel(s), [2] the exact re-rank kernel(s), [3] overflow fallbacks. */
SPHB_API int sphb_knn_last_timing(float* ms4);
/* Global ids of the query rows of the last kNN call that left the main 1-product sweep: which = 0 rows re-swept with the
 * 3-product filter (their candidate band overflowed), which = 1 rows recomputed by the exact SIMT kernel.  Writes at most
 * cap ids to out; *count receives the total.  Diagnostic (parity checks sample these rows); no reference counterpart. */
SPHB_API int sphb_knn_last_overflow_rows(int which, int64_t* out, int64_t cap, int64_t* count);
/*
 * Stage A followed by the level-0 P rows of Stage B in one call — the pair
 * ComputeHierarchy::compute() runs back to back (computeKnnGraph, sph/ComputeHierarchy.cpp:97-159,
 * then ImageHierarchy::computePreparations -> normalizeKnnDistances, sph/ImageHierarchy.cpp:153).
 * x goes up once, the graph stays in HBM between the stages, the four outputs come back once.
 * Query rows [q0,q1): ids/dist (q1-q0)*k, cols/vals (q1-q0)*(k-1).
 */
SPHB_API int sphb_knn_graph_prow(const float* x, int64_t n, int d, int k, int metric, int l2_squared, float perplexity,
                                 int64_t q0, int64_t q1, int64_t* ids, float* dist, uint32_t* cols, float* vals);

/*
 * Exact kNN of an explicit list of query rows WITHOUT any device allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  Same result contract as sphb_knn_exact (fp64 pair keys, (distance, index) order,
 * fused post-passes); output row r holds the neighbours of point d_rows[r].  All pointers are DEVICE pointers.
 * The scratch (per-segment candidate lists) is a fixed 8 MiB device array inside the library, held by every GPU that
 * loads the library until exit.  Limits: nrows <= 256 and k <= 128 per call (SPHB_ERR_CAPACITY beyond).  The call makes no
 * cudaMalloc / cudaFree / allocating library call; the first call on a device looks up the array's address once
 * (cudaGetSymbolAddress) and, if sphb_init was not called yet, creates the library stream — make one call before a
 * graph capture.  All calls share the array: calls that overlap on the GPU (two streams, two threads) race.
 */
SPHB_API int sphb_knn_rows_noalloc_dev(const float* d_x, int64_t n, int d, int k, int metric, int l2_squared, const int64_t* d_rows,
                                       int64_t nrows, int64_t* d_ids, float* d_dist, void* stream);

/* Single process, several GPUs: the same result as sphb_knn_graph_prow over the whole stack (q0 = 0, q1 = n), computed on the
 * `num_devices` GPUs listed in `devices` from ONE call of one host thread — the form a drop-in for the single-threaded
 * ComputeHierarchy::compute() (sph/ComputeHierarchy.cpp:262-276) needs on a multi-GPU box.  One worker thread per device:
 * worker g uploads rows [g*ceil(n/G), (g+1)*ceil(n/G)) from `x`, the workers exchange their slabs device to device
 * (cudaMemcpyPeerAsync; NVLink where peer access exists) so that every GPU holds the whole database, each searches its slab of
 * query rows and writes its range of ids / dist / cols / vals.  Host-to-device traffic N*d*4 bytes in total.  Contexts are per
 * host thread: the calling thread's own context (sphb_init, sphb_knn_last_stats, ...) is not touched by this call. */
SPHB_API int sphb_knn_graph_prow_multi(const float* x, int64_t n, int d, int k, int metric, int l2_squared, float perplexity,
                                       const int* devices, int num_devices, int64_t* ids, float* dist, uint32_t* cols, float* vals);

/* ---- Stage B: perplexity-calibrated probability rows -------------------- */
/*
 * Replaces utils::computeGaussianDistributions(const GraphBaseInterface&, vector<T>&, float)
 *   (sph/utils/GraphNormalization.hpp:16-21, .cpp:191-338) on a fixed-k graph whose
 * first neighbour is the point itself.  perplexity <= 0 means (k-1)/3.
 * cols/vals: n*(k-1), each row column-sorted (MapMemEff / SparseVector order).
 * sigma (optional, may be NULL): n doubles.
 */
SPHB_API int sphb_prow_gaussian_graph(const int64_t* ids, const float* dist, int64_t n, int k, float perplexity,
                                      uint32_t* cols, float* vals, double* sigma);
SPHB_API int sphb_prow_gaussian_graph_dev(const int64_t* d_ids, const float* d_dist, int64_t n, int k, float perplexity,
                                          uint32_t* d_cols, float* d_vals, double* d_sigma, void* stream);
/* The same two row builders for a RAGGED graph (utils::Graph with per-point neighbour counts — merged or symmetrised
 * graphs; both reference functions take any GraphBaseInterface, sph/utils/GraphNormalization.cpp:191-411; reached with
 * the merged data graph from Hierarchy::updateRandomWalks, sph/utils/Hierarchy.cpp:317-336).  rowptr int64[n+1], every row
 * starts with the point itself; linear = 0 Gaussian (perplexity <= 0 means (k_p - 1) / 3 per row), 1 linear.
 * out_rowptr[n+1] (row p holds k_p - 1 column-sorted entries), out_cols / out_vals: rowptr[n] - n entries. */
SPHB_API int sphb_prow_graph_ragged(int64_t n, const int64_t* rowptr, const int64_t* ids, const float* dist, float perplexity, int linear,
                                    int64_t* out_rowptr, uint32_t* out_cols, float* out_vals);
/* utils::computeLinearDistributions (sph/utils/GraphNormalization.cpp:340-411). */
SPHB_API int sphb_prow_linear_graph(const int64_t* ids, const float* dist, int64_t n, int k,
                                    uint32_t* cols, float* vals);
/* utils::computeGaussianDistributions(vector<T>&, float, int) in place on sparse rows
 *   (sph/utils/GraphNormalization.cpp:38-189); ignore = -1 on level >= 1. */
SPHB_API int sphb_prow_gaussian_inplace(int64_t n, const int64_t* rowptr, float* val, float perplexity, int ignore);
/* utils::computeGaussianDistributions(dist², idx, nn, rows, params) flat variant
 *   (sph/utils/HDILibHelper.hpp:112-140; tol 1e-5): fills temp[n*k]; the caller scatters. */
SPHB_API int sphb_prow_gaussian_flat(const float* dist, int64_t n, int k, double perplexity, float* temp);
/* SURVEY 8(f)-3 — utils::computeExponentialDistributions (sph/utils/GraphNormalization.hpp:29-32, .cpp:413-593): the
 * UMAP-flavoured rows (NormalizationScheme::UMAP, sph/LevelSimilarities.cpp:499-503,532-533).  The reference passes each
 * row to umappp::neighbor_similarities (libscran-umappp 3.2.0, default options, fp32): rho = first non-zero distance,
 * sigma from sum_j [d_j > rho ? exp(-(d_j - rho)/sigma) : 1] = log2(n_neighbours + 1) (|sum - target| < 1e-5, <= 64
 * evaluations), sigma >= 1e-3 * mean(d), w_j = d_j > rho ? exp(-(d_j - rho)/sigma) : 1.  umappp is not part of the
 * reference tree, so the arithmetic follows the published algorithm; results agree with any conforming implementation to
 * the search tolerance (parity unpinned).
 *  _graph: (graphView, rows) overload (.cpp:508-593) on a fixed-k graph, neighbour 0 (the point itself) skipped; cols/vals:
 *          n*(k-1); sort_cols = 0 keeps the neighbour order (T = MapMemEff, storage.emplace_back .cpp:566-569), 1 sorts each
 *          row by column (T = SparseVecSPH, insert .cpp:559-562).
 *  _inplace: (rows, ignore) overload (.cpp:413-506) on CSR rows, values overwritten; an entry whose COLUMN equals `ignore`
 *          (>= 0; the reference calls it with -1) is left out of the row and keeps its value; col may be NULL when ignore < 0.
 *          Rows of at most 1024 entries. */
SPHB_API int sphb_prow_exponential_graph(const int64_t* ids, const float* dist, int64_t n, int k, int sort_cols,
                                         uint32_t* cols, float* vals);
SPHB_API int sphb_prow_exponential_graph_dev(const int64_t* d_ids, const float* d_dist, int64_t n, int k, int sort_cols,
                                             uint32_t* d_cols, float* d_vals, void* stream);
SPHB_API int sphb_prow_exponential_inplace(int64_t n, const int64_t* rowptr, const uint32_t* col, float* val, int ignore);
SPHB_API int sphb_prow_exponential_inplace_dev(int64_t n, const int64_t* d_rowptr, const uint32_t* d_col, float* d_val,
                                               int max_row_nnz, int ignore, void* stream);
/* utils::symmetrizeTSNE / symmetrizeUMAP (sph/utils/HDILibHelper.hpp:260-302).
 * out_* capacity `cap` entries (2*nnz always suffices); *out_nnz receives the size. */
SPHB_API int sphb_symmetrize(int64_t n, const int64_t* rowptr, const uint32_t* col, const float* val, int umap,
                             int64_t* out_rowptr, uint32_t* out_col, float* out_val, int64_t cap, int64_t* out_nnz);

/* ---- Stage C: random walks --------------------------------------------- */
/*
 * Replaces utils::doRandomWalks (sph/utils/SparseMatrixAlgorithms.hpp:73, .cpp:34-290).
 * Counter-based RNG (Philox4x32-10 keyed by seed, start point, walk, step): the
 * result is independent of thread layout; parity with the reference's
 * mt19937_64 streams is statistical.
 */
SPHB_API int sphb_random_walks(int64_t n, const int64_t* rowptr, const uint32_t* col, const float* val,
                               uint64_t num_walks, uint64_t walk_len, int weighting, float prune_value, uint64_t prune_steps,
                               uint64_t seed, int remove_diagonal, int normalize,
                               int64_t* out_rowptr, uint32_t* out_col, float* out_val, int64_t cap, int64_t* out_nnz);
/* The same for start points [q0, q1) only — the multi-GPU shard of doRandomWalks (start points are independent, the
 * transition matrix is replicated; the RNG is keyed by the global start id, so slabs concatenate to the full result).
 * out_rowptr: q1 - q0 + 1 entries relative to the slab. */
SPHB_API int sphb_random_walks_slab(int64_t n, const int64_t* rowptr, const uint32_t* col, const float* val,
                                    uint64_t num_walks, uint64_t walk_len, int weighting, float prune_value, uint64_t prune_steps,
                                    uint64_t seed, int remove_diagonal, int normalize, int64_t q0, int64_t q1,
                                    int64_t* out_rowptr, uint32_t* out_col, float* out_val, int64_t cap, int64_t* out_nnz);
/* utils::mergeNodesRandomWalks (sph/utils/SparseMatrixAlgorithms.hpp:82, .cpp:292-352). */
SPHB_API int sphb_rw_merge(int64_t n, const int64_t* rowptr, const uint32_t* col, const float* val,
                           const uint64_t* parents, int64_t n_merged, int norm, int weight_by_size,
                           int64_t* out_rowptr, uint32_t* out_col, float* out_val, int64_t cap, int64_t* out_nnz);

/* ---- Stage D: level similarities ---------------------------------------- */
/*
 * Replaces utils::createSimilarities (sph/utils/SparseMatrixAlgorithms.hpp:121, .cpp:963-1378):
 * -log(sqrt(A)·sqrt(A)ᵀ), pruned, per-row bottom-k by (value, column), column-sorted,
 * L1-normalised.  variant: 0 auto (n > 5000 -> HDI path), 1 HDI, 2 Eigen.
 * comp_sizes may be NULL.  out: fixed stride k, cols padded 0xFFFFFFFF, row_nnz[n].
 */
SPHB_API int sphb_create_similarities(int64_t n, const int64_t* rowptr, const uint32_t* col, const float* val,
                                      int64_t k, float prune_val, int variant, const int64_t* comp_sizes,
                                      uint32_t* out_cols, float* out_vals, int32_t* row_nnz);
/* Batched utils::randomWalksBhattacharyya (mode 0) / single overlap (mode 1)
 *   (sph/utils/Similarities.cpp:374-396) over an edge list. */
SPHB_API int sphb_pair_sims(int64_t n, const int64_t* rowptr, const uint32_t* col, const float* val,
                            const int64_t* pa, const int64_t* pb, int64_t npairs, int mode, float* out);

/* Batched utils::findTopK (top = 1) / findBottomK (top = 0) (sph/utils/SparseMatrixAlgorithms.cpp:720-776, deque
 * admission rules of sph/utils/MaxSizeDeque.hpp:97-200) over the rows of a CSR matrix — the per-row conversion
 * convertEigenSparseVecToHDILibSparseVec does when randomWalkPairSims = false (sph/LevelSimilarities.cpp:481-486).
 * out: fixed stride k, column-sorted, cols padded 0xFFFFFFFF, row_nnz[n]. */
SPHB_API int sphb_find_k(int64_t n, const int64_t* rowptr, const uint32_t* col, const float* val, int64_t k, int top,
                         uint32_t* out_cols, float* out_vals, int32_t* row_nnz);
/* Batched utils::simNeighborOverlap (sph/utils/Similarities.cpp:174-228): for each pair of components (pa, pb)
 * card(U_a ∩ U_b) / min(card U_a, card U_b), U_c = sorted-unique union of the data-level kNN lists (knn_ids [n*k]) of
 * the pixels comp_pix[comp_ptr[c] .. comp_ptr[c+1]) of component c (Hierarchy::getRepresentedDataPoints).
 * componentDistance = 1 - sim (sph/utils/Similarities.cpp:123-156). */
SPHB_API int sphb_neighbor_overlap(const int64_t* knn_ids, int64_t n, int k, int64_t ncomp, const int64_t* comp_ptr,
                                   const int64_t* comp_pix, const int64_t* pa, const int64_t* pb, int64_t npairs, float* out);

/* ---- ragged-graph passes next to the kNN search and between levels -------- */
/*
 * A graph crosses the boundary as utils::Graph stores it (sph/utils/Graph.hpp:155-273): rowptr int64[n+1]
 * (offsets + total), ids int64[nnz] (knnIndices), dist float[nnz] (knnDistances).  rowptr == NULL means a
 * fixed-k graph (Graph::updateFixedNumNeighbors): row r is [r*k, (r+1)*k).
 *
 * Replaces utils::symmetrizeGraph (sph/utils/GraphUtils.cpp:197-317), the step NearestNeighbors::compute runs
 * right after the search when symmetricNeighbors = true (sph/NearestNeighbors.cpp:176-183): for every stored
 * (i, j) or (j, i) both rows get the entry, distance min(d_ij, d_ji) where both exist (first stored occurrence of
 * each), rows ordered by (distance, index), then ensureClosestPointIsSelf (sph/utils/GraphUtils.cpp:23-95).
 * Output capacity `cap` entries (2*nnz always suffices); *out_nnz receives the size.
 */
SPHB_API int sphb_symmetrize_graph(int64_t n, const int64_t* rowptr, int k, const int64_t* ids, const float* dist,
                                   int64_t* out_rowptr, int64_t* out_ids, float* out_dist, int64_t cap, int64_t* out_nnz);
/* Connected components of the kNN graph with edges taken as undirected — utils::labelGraphWeakComponents, and
 * labelGraphStrongComponents on a symmetric graph (sph/utils/GraphUtils.cpp:371-423), as
 * NearestNeighbors::computeConnectedComponents uses them (sph/NearestNeighbors.cpp:318-409).  labels[n]: component of
 * every point, components numbered in ascending order of their lowest point (the order Boost reaches them);
 * *num_components receives their number. */
SPHB_API int sphb_weak_components(int64_t n, const int64_t* rowptr, int k, const int64_t* ids, int64_t* labels, int64_t* num_components);
/* SURVEY 8(f)-2 — component bridging, the device parts of NearestNeighbors::connectComponents
 * (sph/NearestNeighbors.cpp:494-861; forced by GEO_CENTROID configurations, sph/ComputeHierarchy.cpp:41-42).
 * labels[n] in [0, num_components) as sphb_weak_components writes them.
 *  sphb_component_centers: centre of mass of every component (computeCentersOfMass :655-680 with utils::computeMean,
 *    sph/utils/Math.hpp:29-48): fp64 sums in ascending point order, divided by the count, rounded to fp32 — bit-identical
 *    to the reference's sequential sums.  centers: num_components * d.
 *  sphb_bridge_components: for every edge (edges[2e] = component A, edges[2e+1] = component B — the source and target of
 *    an edge of the minimum spanning tree the host builds over the centres, :682-708) the pair the reference connects
 *    (insertConnectionsBetweenComponents :592-653): for every point of B its nearest point of A, then the closest such
 *    pair; ids_a[e] / ids_b[e] = global point ids, dist[e] = their SQUARED L2 distance (what faiss::IndexFlat returns and
 *    the reference inserts into the graph as is).  Exact: ties resolve to the lowest point of B, then of A.  The reference
 *    switches to the approximate IndexIVFFlat (nlist 100, nprobe 10) when A holds more than 3900 points (:607-626); this
 *    entry point is exact at every size, so its distance is <= the reference's there.  Only SPHB_METRIC_L2.
 *    The host keeps the MST (Kruskal over num_components^2 / 2 centre distances) and the row insertions (:547-572);
 *    sph_b200/host.py NearestNeighbors.connectComponents is that mirror.  `edges` is always a HOST array. */
SPHB_API int sphb_component_centers(const float* x, int64_t n, int d, const int64_t* labels, int64_t num_components, float* centers);
SPHB_API int sphb_component_centers_dev(const float* d_x, int64_t n, int d, const int64_t* d_labels, int64_t num_components,
                                        float* d_centers, void* stream);
SPHB_API int sphb_bridge_components(const float* x, int64_t n, int d, int metric, const int64_t* labels, int64_t num_components,
                                    const int64_t* edges, int64_t num_edges, int64_t* ids_a, int64_t* ids_b, float* dist);
SPHB_API int sphb_bridge_components_dev(const float* d_x, int64_t n, int d, int metric, const int64_t* d_labels,
                                        int64_t num_components, const int64_t* edges, int64_t num_edges, int64_t* d_ids_a,
                                        int64_t* d_ids_b, float* d_dist, void* stream);
/* Device-pointer variant (graph resident in HBM, e.g. straight from sphb_knn_exact_dev); d_rowptr may be NULL for a
 * fixed-k graph, nnz = number of stored entries; *out_nnz is written on the host (one stream synchronisation). */
SPHB_API int sphb_symmetrize_graph_dev(int64_t n, const int64_t* d_rowptr, int k, const int64_t* d_ids, const float* d_dist, int64_t nnz,
                                       int64_t* d_out_rowptr, int64_t* d_out_ids, float* d_out_dist, int64_t cap, int64_t* out_nnz,
                                       void* stream);
/* utils::mergeGraphNodes (sph/utils/SparseMatrixAlgorithms.cpp:486-561): node r merges into parents[r]; the
 * distance of (parent(r), parent(c)) is the minimum over the children; rows ordered by (distance, index), self first.
 * cap: nnz of the input always suffices. */
SPHB_API int sphb_merge_graph_nodes(int64_t n, const int64_t* rowptr, int k, const int64_t* ids, const float* dist,
                                    const uint64_t* parents, int64_t n_merged,
                                    int64_t* out_rowptr, int64_t* out_ids, float* out_dist, int64_t cap, int64_t* out_nnz);
/* utils::mergeNodesDataDistances (sph/utils/SparseMatrixAlgorithms.cpp:443-484), the sparse-row form of the same
 * merge: a stored or missing 0 counts as empty and is replaced by the next value, otherwise the smaller value stays. */
SPHB_API int sphb_merge_data_distances(int64_t n, const int64_t* rowptr, const uint32_t* col, const float* val,
                                       const uint64_t* parents, int64_t n_merged,
                                       int64_t* out_rowptr, uint32_t* out_col, float* out_val, int64_t cap, int64_t* out_nnz);
/* useRandomWalksAsKnnDistances (sph/LevelSimilarities.cpp:346-389): every non-zero random-walk entry becomes a
 * neighbour at distance 1 - value; rows stable-sorted by distance (ties keep column order).  out_ids/out_dist: nnz. */
SPHB_API int sphb_rw_as_knn_distances(int64_t n, const int64_t* rowptr, const uint32_t* col, const float* val,
                                      int64_t* out_rowptr, int64_t* out_ids, float* out_dist, int64_t* out_nnz);
/* Zero removal of the level rows (sph/LevelSimilarities.cpp:566-581).  out_col/out_val: nnz. */
SPHB_API int sphb_remove_zeros(int64_t n, const int64_t* rowptr, const uint32_t* col, const float* val,
                               int64_t* out_rowptr, uint32_t* out_col, float* out_val, int64_t* out_nnz);
/* The post-merge step of Hierarchy::updateRandomWalks (sph/utils/Hierarchy.cpp:287-307) on the merged walks:
 * remove_diagonal != 0: removeDiagonalElements(keepSingleEntry = true) (sph/utils/SparseMatrixAlgorithms.cpp:704-718) — a row with more
 * than one entry loses its diagonal and every stored exact zero; norm: 0 none, 1 NormType::ONEDIM (every row / its fp32 sum,
 * normalizeUnitSparseVector :666-673), 2 NormType::TWODIM (every value / the fp32 sum over all rows, normalizeSparseMatrix :617-628).
 * out_col / out_val: nnz of the input suffices. */
SPHB_API int sphb_rw_post_merge(int64_t n, const int64_t* rowptr, const uint32_t* col, const float* val, int remove_diagonal, int norm,
                                int64_t* out_rowptr, uint32_t* out_col, float* out_val, int64_t* out_nnz);
SPHB_API int sphb_rw_post_merge_dev(int64_t n, const int64_t* d_rowptr, const uint32_t* d_col, const float* d_val, int remove_diagonal, int norm,
                                    int64_t* d_out_rowptr, uint32_t* d_out_col, float* d_out_val, int64_t* out_nnz, void* stream);
/* Exact level kNN in a random-walk component metric — LevelSimilarities' computeExactKnn
 * (sph/LevelSimilarities.cpp:211-252) in NeighborWalksBhattacharyyaSpace (mode 0,
 * sph/utils/NeighborWalksBhattacharyyaSpace.hpp:28-41: distance 1 - randomWalksBhattacharyya(i, j)) or
 * NeighborWalksSingleOverlapSpace (mode 1, sph/utils/NeighborWalksSingleOverlapSpace.hpp:26-36: distance 1 - A[i][j]);
 * like the reference's spaces, j == i is not special-cased.  (rowptr, col, val) is the level's random-walk matrix
 * (n x n CSR).  out: n*k, each row ordered by (distance, index) — the reference's unstable sort leaves ties unordered;
 * slots beyond n keep -1 / -1.0f. */
SPHB_API int sphb_level_knn_walks(int64_t n, const int64_t* rowptr, const uint32_t* col, const float* val, int mode, int64_t k,
                                  int64_t* out_ids, float* out_dist);
/* The same exact level kNN in the Euclid metric (EuclidDistSpace, sph/utils/EuclidDistSpace.hpp:29-39; computeExactKnn
 * sph/LevelSimilarities.cpp:211-252 as reached from :414): the symmetric Hausdorff distance (sphb_euclid_hausdorff) of every
 * component against every component, the k smallest (distance, index) of every row selected on the device.
 * out: ncomp*k, rows ordered by (distance, index); k <= ncomp. */
SPHB_API int sphb_level_knn_euclid(const float* x, int64_t n, int d, int64_t ncomp, const int64_t* comp_ptr, const int64_t* comp_pix, int64_t k,
                                   int64_t* out_ids, float* out_dist);
/* SURVEY 8(f)-4 — geodesic component distances, the metric of GEO_CENTROID / GEO_WALKS: utils::geodesicDistance
 * (sph/utils/Similarities.cpp:234-309) over utils::computeShortestPath (sph/utils/ShortestPath.cpp:100-171), batched.
 * The graph is the data-level kNN graph the reference hands to createBoostGraph (sph/utils/AStarBoost.cpp:104-133; with
 * GEO_* the connected graph, sph/ComputeHierarchy.cpp:135): rowptr int64[n+1] or NULL for a fixed-k graph (ids / dist n*k),
 * one directed edge per stored (point, neighbour != point, distance); distances must be non-negative.  labels (may be NULL):
 * connected-component label of every point (sph/ComputeHierarchy.cpp:143) — points with different labels have no path.
 * Per (start, end): different labels -> no path; end stored in start's row -> the stored distance of the first such slot;
 * else the shortest-path distance, i.e. the minimum over the paths of the left-to-right fp32 sum of the edge weights (what
 * dijkstraBoost settles, AStarBoost.cpp:241-257, and what astarBoost returns with an admissible heuristic — true L2 edge
 * weights, L2squared = false).  One single-source search per distinct start point of a call, on the device.
 *
 * sphb_geodesic_distances: computeShortestPath for npairs point pairs, -1 where there is no path.
 * sphb_geodesic_hausdorff: geodesicDistance for npairs component pairs (pa[p], pb[p]), components as in
 * sphb_euclid_hausdorff: symmetric Hausdorff distance (sph/utils/Math.cpp:167-172) of the matrix of computeShortestPath
 * values from the points of pa[p] to the points of pb[p], "no path" entering as FLT_MAX (:278-281); single-point components
 * give the level-0 branch (:245-252).  The sub-sampled branch (:288-303) draws from an unseeded random_device: the caller
 * passes the sampled lists.  sphb_geodesic_last_stats: {vertices expanded, bitmap passes, searches run, sources} of the
 * calling thread's last call. */
SPHB_API int sphb_geodesic_distances(int64_t n, const int64_t* rowptr, int k, const int64_t* ids, const float* dist, const int64_t* labels,
                                     const int64_t* start, const int64_t* end, int64_t npairs, float* out);
SPHB_API int sphb_geodesic_hausdorff(int64_t n, const int64_t* rowptr, int k, const int64_t* ids, const float* dist, const int64_t* labels,
                                     int64_t ncomp, const int64_t* comp_ptr, const int64_t* comp_pix, const int64_t* pa, const int64_t* pb,
                                     int64_t npairs, float* out);
SPHB_API int sphb_geodesic_last_stats(int64_t* out4);
/* The same two calls on a graph that stays on the device between calls: ImageHierarchy::computeDistances asks for the distances of
 * every level on the same data graph (sph/ImageHierarchy.cpp:192-249), and uploading it is about half of a level-0 call.  What
 * _hierarchy.bgraph is to the reference (built once, sph/ImageHierarchy.cpp:427-431).  A handle belongs to the thread (device
 * context) that created it; destroy it there. */
SPHB_API int sphb_geodesic_graph_create(int64_t n, const int64_t* rowptr, int k, const int64_t* ids, const float* dist, const int64_t* labels,
                                        void** handle);
SPHB_API int sphb_geodesic_graph_destroy(void* handle);
SPHB_API int sphb_geodesic_distances_g(void* handle, const int64_t* start, const int64_t* end, int64_t npairs, float* out);
SPHB_API int sphb_geodesic_hausdorff_g(void* handle, int64_t ncomp, const int64_t* comp_ptr, const int64_t* comp_pix, const int64_t* pa,
                                       const int64_t* pb, int64_t npairs, float* out);
/* SURVEY 8(f)-4, cache blobs — the reference's stage-cache files (sph/utils/FileIO.hpp:165-291, 523-640; FileIO.cpp:586-618,
 * 620-799, 895-937, 1274-1460), written and read on the HOST so that results of this back-end can be dropped into the
 * reference's cache directory and loaded by it, and the reverse.  Blocks are LZ4 block format (LZ4_compress_default /
 * LZ4_decompress_safe); files above INT_MAX bytes use the reference's chunked layout.
 *   vec:   writeCompressedVecToBinary<T> / loadCompressedVecFromBinary<T> (connected-component labels, spanning-tree ids, ...)
 *   graph: writeCompressedGraphToBinary / loadCompressedGraphFromBinary — <base>Distances.cache, <base>Indices.cache,
 *          <base>NNs.cache, <base>Symmetric.cache (kNN / symmetric / connected graphs, sph/NearestNeighbors.cpp:1011-1017; level
 *          graphs, sph/LevelSimilarities.cpp:744); rowptr NULL = fixed k
 *   prob:  writeCompressedSparseMatHDIToBinary / loadCompressedSparseMatHDIFromBinary — a P matrix as CSR
 *          (_lls_probDist<level>.cache, sph/LevelSimilarities.cpp:763-787)
 * Readers called with NULL output arrays fill only the sizes (*count / *n, *nnz). */
SPHB_API int64_t sphb_lz4_compress_bound(int64_t n);
SPHB_API int64_t sphb_lz4_compress(const void* src, int64_t n, void* dst, int64_t cap);   /* bytes written, -1 on failure */
SPHB_API int64_t sphb_lz4_decompress(const void* src, int64_t n, void* dst, int64_t cap); /* bytes written, -1 on a malformed block */
SPHB_API int sphb_cache_write_vec(const char* path, const void* data, int64_t count, int elem_size);
SPHB_API int sphb_cache_read_vec(const char* path, void* out, int64_t cap, int elem_size, int64_t* count);
SPHB_API int sphb_cache_write_graph(const char* base, int64_t n, const int64_t* rowptr, int k, const int64_t* ids, const float* dist, int symmetric);
SPHB_API int sphb_cache_read_graph(const char* base, int64_t cap_n, int64_t cap_nnz, int64_t* n, int64_t* nnz, int64_t* nns, int64_t* ids,
                                   float* dist, int* symmetric);
SPHB_API int sphb_cache_write_prob(const char* path, int64_t n, const int64_t* rowptr, const uint32_t* col, const float* val);
SPHB_API int sphb_cache_read_prob(const char* path, int64_t cap_n, int64_t cap_nnz, int64_t* n, int64_t* nnz, int64_t* rowptr, uint32_t* col,
                                  float* val);
/* The same exact level kNN in the neighbour-overlap metric (NeighborOverlapSpace, sph/utils/NeighborOverlapSpace.hpp:31-43;
 * the default componentSim of LevelSimilaritiesSettings, sph/utils/Settings.hpp:170): distance 1 - simNeighborOverlap
 * (sph/utils/Similarities.cpp:174-228) over the components' kNN-id unions; arguments as sphb_neighbor_overlap.
 * out: ncomp*k, rows ordered by (distance, index). */
SPHB_API int sphb_level_knn_overlap(const int64_t* knn_ids, int64_t n, int k_data, int64_t ncomp, const int64_t* comp_ptr, const int64_t* comp_pix,
                                    int64_t k, int64_t* out_ids, float* out_dist);
/* Batched utils::euclidDistance (sph/utils/Similarities.cpp:414-483): symmetric Hausdorff distance
 * (sph/utils/Math.cpp:167-172) of the all-pairs L2 matrix between the pixels of components pa[p] and pb[p]
 * (comp_pix[comp_ptr[c] .. comp_ptr[c+1]), as Hierarchy::getRepresentedDataPoints lists them; single-pixel
 * components give the level-0 branch).  The reference sub-samples components larger than numGeodesicSamples with an
 * unseeded random_device (:455-474); the caller passes the sampled lists in that case. */
SPHB_API int sphb_euclid_hausdorff(const float* x, int64_t n, int d, int64_t ncomp, const int64_t* comp_ptr, const int64_t* comp_pix,
                                   const int64_t* pa, const int64_t* pb, int64_t npairs, float* out);

/* ---- resident level pipeline: device-pointer variants + the fused level call ---------------
 * Same semantics as the host-pointer entry points above; all pointers prefixed d_ are DEVICE pointers, `stream` is a
 * cudaStream_t, counts (*out_nnz) are written on the host after one stream synchronisation.  A caller that keeps the
 * matrices of LevelSimilarities::compute (sph/LevelSimilarities.cpp:118-182) in HBM binds these and moves nothing over
 * PCIe between the stages. */
/* utils::doRandomWalks for start points [q0, q1) (sph/utils/SparseMatrixAlgorithms.cpp:34-290). */
SPHB_API int sphb_random_walks_dev(int64_t n, const int64_t* d_rowptr, const uint32_t* d_col, const float* d_val, uint64_t num_walks,
                                   uint64_t walk_len, int weighting, float prune_value, uint64_t prune_steps, uint64_t seed, int remove_diagonal,
                                   int normalize, int64_t q0, int64_t q1, int64_t* d_out_rowptr, uint32_t* d_out_col, float* d_out_val, int64_t cap,
                                   int64_t* out_nnz, void* stream);
/* utils::mergeNodesRandomWalks (sph/utils/SparseMatrixAlgorithms.cpp:292-352). */
SPHB_API int sphb_rw_merge_dev(int64_t n, const int64_t* d_rowptr, const uint32_t* d_col, const float* d_val, int64_t nnz, const uint64_t* d_parents,
                               int64_t n_merged, int norm, int weight_by_size, int64_t* d_out_rowptr, uint32_t* d_out_col, float* d_out_val,
                               int64_t cap, int64_t* out_nnz, void* stream);
/* utils::createSimilarities (sph/utils/SparseMatrixAlgorithms.cpp:963-1378); outputs as sphb_create_similarities. */
SPHB_API int sphb_create_similarities_dev(int64_t n, const int64_t* d_rowptr, const uint32_t* d_col, const float* d_val, int64_t nnz, int64_t k,
                                          float prune_val, int variant, const int64_t* d_comp_sizes, uint32_t* d_out_cols, float* d_out_vals,
                                          int32_t* d_row_nnz, void* stream);
/* utils::computeGaussianDistributions in place (sph/utils/GraphNormalization.cpp:38-189); max_row_nnz <= 0: computed on the device. */
SPHB_API int sphb_prow_gaussian_inplace_dev(int64_t n, const int64_t* d_rowptr, float* d_val, int max_row_nnz, float perplexity, int ignore,
                                            void* stream);
/* utils::symmetrizeTSNE / symmetrizeUMAP (sph/utils/HDILibHelper.hpp:260-302). */
SPHB_API int sphb_symmetrize_dev(int64_t n, const int64_t* d_rowptr, const uint32_t* d_col, const float* d_val, int64_t nnz, int umap,
                                 int64_t* d_out_rowptr, uint32_t* d_out_col, float* d_out_val, int64_t cap, int64_t* out_nnz, void* stream);
/* utils::findTopK / findBottomK batched (sph/utils/SparseMatrixAlgorithms.cpp:720-776); max_row_nnz <= 0: computed on the device. */
SPHB_API int sphb_find_k_dev(int64_t n, const int64_t* d_rowptr, const uint32_t* d_col, const float* d_val, int max_row_nnz, int64_t k, int top,
                             uint32_t* d_out_cols, float* d_out_vals, int32_t* d_row_nnz, void* stream);
/* SURVEY 8(e), createSimilarities sharded by OUTPUT ROW: rows [r0, r1) of the result only (every output row needs the whole
 * random-walk matrix and nothing of the other output rows, so slabs computed on different devices concatenate to exactly the
 * one-call result).  Outputs hold r1 - r0 rows in the layout of sphb_create_similarities. */
SPHB_API int sphb_create_similarities_slab(int64_t n, const int64_t* rowptr, const uint32_t* col, const float* val, int64_t k, float prune_val,
                                           int variant, const int64_t* comp_sizes, int64_t r0, int64_t r1, uint32_t* out_cols, float* out_vals,
                                           int32_t* row_nnz);
SPHB_API int sphb_create_similarities_slab_dev(int64_t n, const int64_t* d_rowptr, const uint32_t* d_col, const float* d_val, int64_t nnz, int64_t k,
                                               float prune_val, int variant, const int64_t* d_comp_sizes, int64_t r0, int64_t r1,
                                               uint32_t* d_out_cols, float* d_out_vals, int32_t* d_row_nnz, void* stream);
/* The part of sphb_level_prob_dist_dev behind createSimilarities / the top-k conversion: fixed-stride rows (n x k, cols padded
 * 0xFFFFFFFF, row_nnz[n] — e.g. the gathered slabs of sphb_create_similarities_slab_dev) -> CSR -> Gaussian rows in place
 * (scheme 1) -> zero removal -> symmetrizeTSNE (symmetrize != 0).  Output as sphb_level_prob_dist_dev. */
SPHB_API int sphb_level_from_sims_dev(int64_t n, int64_t k, const uint32_t* d_cols, const float* d_vals, const int32_t* d_row_nnz, float perplexity,
                                      int scheme, int symmetrize, int64_t* d_out_rowptr, uint32_t* d_out_col, float* d_out_val, int64_t cap,
                                      int64_t* out_nnz, void* stream);
/*
 * One level of LevelSimilarities::computeProbDistOnLevel's useRandomWalks branch + symmetrizeOutput in ONE call on the level's
 * random-walk matrix (sph/LevelSimilarities.cpp:460-508, 566-581, 589-623), every intermediate resident in HBM:
 *   pair_sims = 1: createSimilarities(walks, k, prune_val, comp_sizes) (randomWalkPairSims, the default),
 *   pair_sims = 0: per-row top-k conversion (convertEigenSparseVecToHDILibSparseVec);
 *   scheme 1 (NormalizationScheme::TSNE): computeGaussianDistributions(rows, perplexity, ignore = -1) in place; 0: none;
 *   exact zeros erased; symmetrize != 0 (with scheme 1): symmetrizeTSNE.
 * Output CSR (the level's SparseMatHDI): out_rowptr[n+1], out_col / out_val with capacity `cap` entries
 * (2 * n * k always suffices); *out_nnz receives the size.
 */
SPHB_API int sphb_level_prob_dist_dev(int64_t n, const int64_t* d_rowptr, const uint32_t* d_col, const float* d_val, int64_t nnz, int64_t k,
                                      float prune_val, int variant, const int64_t* d_comp_sizes, int pair_sims, float perplexity, int scheme,
                                      int symmetrize, int64_t* d_out_rowptr, uint32_t* d_out_col, float* d_out_val, int64_t cap, int64_t* out_nnz,
                                      void* stream);
SPHB_API int sphb_level_prob_dist(int64_t n, const int64_t* rowptr, const uint32_t* col, const float* val, int64_t k, float prune_val, int variant,
                                  const int64_t* comp_sizes, int pair_sims, float perplexity, int scheme, int symmetrize, int64_t* out_rowptr,
                                  uint32_t* out_col, float* out_val, int64_t cap, int64_t* out_nnz);
/* device time in ms of the stages of the last sphb_level_prob_dist[_dev] call (CUDA events on its stream):
 * [0] similarities / top-k, [1] compaction to CSR, [2] Gaussian rows, [3] zero removal, [4] symmetrise. */
SPHB_API int sphb_level_last_timing(float* ms5);

#ifdef __cplusplus
}
#endif
#endif /* SPHB_H */
