"""One process per GPU: Stage A + level-0 P rows with the query rows sharded over the ranks.

The database is replicated through NCCL, not through PCIe: every rank uploads only ITS N/G rows of the image stack from
(pinned) host memory and one all-gather over NVLink assembles the full stack on every GPU ("each rank H2D-copies its own
1/G and ncclAllGather"), so the host-to-device traffic of a call is N*d*4 bytes in total, independent
of the number of GPUs.  Each rank then searches its slab of query rows against the full database
(sphb_knn_exact_dev), builds its P rows (sphb_prow_gaussian_graph_dev) and copies its four result slabs into its
range of the caller's host buffers — no collective on the result side.

torch is used for device buffers, streams and torch.distributed (NCCL) only; the kernels are the C-ABI library's.
"""
from __future__ import annotations

import numpy as np

from . import capi


def slab_bounds(n: int, rank: int, world: int) -> tuple[int, int, int]:
    """(q0, q1, chunk): rank r owns rows [r*chunk, min(n, (r+1)*chunk)), chunk = ceil(n / world)."""
    chunk = (n + world - 1) // world
    q0 = min(n, rank * chunk)
    return q0, min(n, q0 + chunk), chunk


class ShardedKnnProw:
    """Reusable buffers for repeated calls on stacks of one shape (n, d, k)."""

    def __init__(self, n: int, d: int, k: int, rank: int, world: int, group=None):
        import torch

        self.n, self.d, self.k, self.rank, self.world, self.group = n, d, k, rank, world, group
        self.q0, self.q1, self.chunk = slab_bounds(n, rank, world)
        nq = self.q1 - self.q0
        dev = torch.device("cuda", torch.cuda.current_device())
        self.x_all = torch.empty((world * self.chunk, d), dtype=torch.float32, device=dev)  # the replicated database (rows >= n: padding)
        self.x_mine = self.x_all[rank * self.chunk:(rank + 1) * self.chunk]
        self.ids = torch.empty((nq, k), dtype=torch.int64, device=dev)
        self.dist = torch.empty((nq, k), dtype=torch.float32, device=dev)
        self.cols = torch.empty((nq, k - 1), dtype=torch.int32, device=dev)
        self.vals = torch.empty((nq, k - 1), dtype=torch.float32, device=dev)

    def replicate(self, x_host) -> None:
        """x_host: torch CPU tensor [n, d] (pinned for full PCIe rate).  Uploads this rank's rows, all-gathers the rest."""
        import torch
        import torch.distributed as dist

        nq = self.q1 - self.q0
        if nq > 0:
            self.x_mine[:nq].copy_(x_host[self.q0:self.q1], non_blocking=True)
        if nq < self.chunk:
            self.x_mine[nq:].zero_()
        if self.world > 1:
            dist.all_gather_into_tensor(self.x_all, self.x_mine.clone(), group=self.group)  # clone: input must not alias the output
        _ = torch

    def compute(self, perplexity: float = 0.0, metric: int = capi.METRIC_L2, l2_squared: bool = False) -> None:
        import torch

        s = torch.cuda.current_stream().cuda_stream
        nq = self.q1 - self.q0
        if nq == 0:
            return
        capi.knn_exact_dev(self.x_all.data_ptr(), self.n, self.d, self.k, metric, l2_squared, self.q0, self.q1, self.ids.data_ptr(),
                           self.dist.data_ptr(), s)
        capi.prow_gaussian_graph_dev(self.ids.data_ptr(), self.dist.data_ptr(), nq, self.k, perplexity, self.cols.data_ptr(), self.vals.data_ptr(), s)

    def download(self, out) -> None:
        """out = (ids, dist, cols, vals): torch CPU tensors for THIS rank's slab (pinned for full rate), e.g. views of shared buffers."""
        for dst, src in zip(out, (self.ids, self.dist, self.cols, self.vals)):
            dst.copy_(src, non_blocking=True)

    def __call__(self, x_host, out, perplexity: float = 0.0, metric: int = capi.METRIC_L2, l2_squared: bool = False) -> None:
        """replicate -> search -> P rows -> download; the graph (ids, dist) goes back to the host on a side stream while the P-row
        kernel runs, as in sphb_knn_graph_prow."""
        import torch

        self.replicate(x_host)
        nq = self.q1 - self.q0
        if nq == 0:
            return
        main = torch.cuda.current_stream()
        s = main.cuda_stream
        capi.knn_exact_dev(self.x_all.data_ptr(), self.n, self.d, self.k, metric, l2_squared, self.q0, self.q1, self.ids.data_ptr(), self.dist.data_ptr(), s)
        if getattr(self, "_side", None) is None:
            self._side = torch.cuda.Stream()
        graph_done = torch.cuda.Event()
        graph_done.record(main)
        self._side.wait_event(graph_done)
        with torch.cuda.stream(self._side):
            out[0].copy_(self.ids, non_blocking=True)
            out[1].copy_(self.dist, non_blocking=True)
        capi.prow_gaussian_graph_dev(self.ids.data_ptr(), self.dist.data_ptr(), nq, self.k, perplexity, self.cols.data_ptr(), self.vals.data_ptr(), s)
        out[2].copy_(self.cols, non_blocking=True)
        out[3].copy_(self.vals, non_blocking=True)
        main.wait_stream(self._side)  # a synchronize on the current stream covers all four copies


def knn_graph_prow_sharded(x: np.ndarray, k: int, rank: int, world: int, perplexity: float = 0.0, group=None):
    """Convenience form: numpy in, this rank's (ids, dist, cols, vals) slabs out."""
    import torch

    n, d = x.shape
    job = ShardedKnnProw(n, d, k, rank, world, group)
    job.replicate(torch.from_numpy(x))
    job.compute(perplexity)
    torch.cuda.synchronize()
    return job.ids.cpu().numpy(), job.dist.cpu().numpy(), job.cols.cpu().numpy().view(np.uint32), job.vals.cpu().numpy()


class ShardedLevel:
    """One level of LevelSimilarities with createSimilarities sharded by OUTPUT ROW: the level's random-walk matrix
    is present on every rank (replicate_csr broadcasts it over NCCL), rank r computes rows slab_bounds(n, r, world) of
    -ln(sqrt(A) sqrt(A)^T) with their bottom-k (sphb_create_similarities_slab_dev — 99 % of a level's time), ONE all-gather of the
    fixed-stride slabs (cols, vals, row_nnz: n * k * 8 bytes), then the cheap tail (CSR -> Gaussian rows -> zero removal ->
    symmetrise, sphb_level_from_sims_dev) on every rank, so every rank ends with the level's whole P matrix in HBM.  An output row
    needs the whole matrix and nothing of the other output rows: the gathered slabs are bit for bit the one-call result."""

    def __init__(self, rank: int, world: int, group=None):
        self.rank, self.world, self.group = rank, world, group

    def _dev(self):
        """Where the buffers live: the current CUDA device (there is no CPU path; _dev / _stream are separate methods so that a
        subclass can place the buffers differently)."""
        import torch

        return torch.device("cuda", torch.cuda.current_device())

    def _stream(self, dev):
        import torch

        return torch.cuda.current_stream().cuda_stream

    def replicate_csr(self, rowptr, col, val, src: int = 0):
        """Tensors on rank `src` (anything, e.g. None, elsewhere) -> the same CSR on every rank."""
        import torch
        import torch.distributed as dist

        if self.world == 1:
            return rowptr, col, val
        dev = self._dev()
        meta = torch.zeros(2, dtype=torch.int64, device=dev)
        if self.rank == src:
            meta[0], meta[1] = rowptr.numel(), col.numel()
        dist.broadcast(meta, src, group=self.group)
        n1, nnz = int(meta[0]), int(meta[1])
        if self.rank != src:
            rowptr = torch.empty(n1, dtype=torch.int64, device=dev)
            col = torch.empty(nnz, dtype=torch.int32, device=dev)
            val = torch.empty(nnz, dtype=torch.float32, device=dev)
        for t in (rowptr, col, val):
            dist.broadcast(t, src, group=self.group)
        return rowptr, col, val

    def __call__(self, rowptr, col, val, k: int, perplexity: float, prune_val: float = 1e-4, variant: int = 0, comp_sizes=None, symmetrize: bool = True):
        """rowptr int64 [n + 1], col int32 (uint32 bits) [nnz], val float32 [nnz]: CUDA tensors, the same on every rank.
        Returns (out_rowptr, out_col, out_val) CUDA tensors of the level's P matrix."""
        import torch
        import torch.distributed as dist

        n = rowptr.numel() - 1
        nnz = col.numel()
        dev = self._dev()
        s = self._stream(dev)
        r0, r1, chunk = slab_bounds(n, self.rank, self.world)
        rows = self.world * chunk
        cols_all = torch.full((rows, k), -1, dtype=torch.int32, device=dev)  # 0xFFFFFFFF padding
        vals_all = torch.zeros((rows, k), dtype=torch.float32, device=dev)
        nnz_all = torch.zeros(rows, dtype=torch.int32, device=dev)
        lo = self.rank * chunk
        if r1 > r0:
            capi.create_similarities_slab_dev(n, rowptr.data_ptr(), col.data_ptr(), val.data_ptr(), nnz, k, prune_val, variant,
                                              comp_sizes.data_ptr() if comp_sizes is not None else None, r0, r1, cols_all[lo:].data_ptr(),
                                              vals_all[lo:].data_ptr(), nnz_all[lo:].data_ptr(), s)
        if self.world > 1:
            for buf in (cols_all, vals_all, nnz_all):
                dist.all_gather_into_tensor(buf, buf[lo:lo + chunk].clone(), group=self.group)  # clone: input must not alias the output
        cap = max(2 * n * k, 1)
        out_rowptr = torch.empty(n + 1, dtype=torch.int64, device=dev)
        out_col = torch.empty(cap, dtype=torch.int32, device=dev)
        out_val = torch.empty(cap, dtype=torch.float32, device=dev)
        m = capi.level_from_sims_dev(n, k, cols_all.data_ptr(), vals_all.data_ptr(), nnz_all.data_ptr(), perplexity, 1, symmetrize, out_rowptr.data_ptr(),
                                     out_col.data_ptr(), out_val.data_ptr(), cap, s)
        return out_rowptr, out_col[:m], out_val[:m]
