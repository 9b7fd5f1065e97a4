"""Host-side mirror of the hot-path entry points of ``metacells.tools`` on numpy / scipy data
(AnnData is not needed: ``what`` is the matrix itself and results are returned instead of stored):

* ``downsample_cells``            tools/downsample.py:25-118
* ``compute_obs_obs_similarity``  tools/similarity.py:29-117, 178-246 (``abs_pearson`` / ``pearson``)
* ``compute_obs_obs_knn_graph``   tools/knn_graph.py:31-121, 203-283
* ``compute_query_atlas_candidates``  the correlation + ordering front-end of tools/project.py:249-270, 309-331

Defaults are the reference's (metacells/parameters.py).
"""
from __future__ import annotations

import ctypes
from typing import Optional, Tuple

import numpy as np
import scipy.sparse as sp

from . import utilities as ut
from ._lib import call, check, host_ptr, lib, row_stride

__all__ = ["downsample_cells", "compute_obs_obs_similarity", "compute_obs_obs_similarity_top", "row_block_of_rank",
           "compute_obs_obs_knn_graph", "KnnGraph", "compute_query_atlas_candidates", "PROJECT_CANDIDATES_COUNT",
           "choose_seeds", "optimize_partitions"]

# metacells/parameters.py
DOWNSAMPLE_MIN_SAMPLES = 750
DOWNSAMPLE_MIN_CELL_QUANTILE = 0.05
DOWNSAMPLE_MAX_CELL_QUANTILE = 0.5
KNN_BALANCED_RANKS_FACTOR = float(np.sqrt(10))
KNN_INCOMING_DEGREE_FACTOR = 3.0
KNN_OUTGOING_DEGREE_FACTOR = 1.0
KNN_MIN_OUTGOING_DEGREE = 2
PROJECT_CANDIDATES_COUNT = 50
MIN_SEED_SIZE_QUANTILE = 0.85
MAX_SEED_SIZE_QUANTILE = 0.95
COOLDOWN_PASS = 0.02
COOLDOWN_NODE = 0.25


def downsample_cells(umis, *, downsample_min_cell_quantile: float = DOWNSAMPLE_MIN_CELL_QUANTILE,
                     downsample_min_samples: float = DOWNSAMPLE_MIN_SAMPLES,
                     downsample_max_cell_quantile: float = DOWNSAMPLE_MAX_CELL_QUANTILE, random_seed: int = 0):
    """tools/downsample.py:84-99: pick the samples count from the cells' totals, downsample every
    cell to it.  Returns ``(downsampled, samples)``."""
    total_per_cell = np.asarray(umis.sum(axis=1)).reshape(-1)
    samples = int(round(min(max(downsample_min_samples, np.quantile(total_per_cell, downsample_min_cell_quantile)),
                            np.quantile(total_per_cell, downsample_max_cell_quantile))))
    return ut.downsample_matrix(umis, per="row", samples=samples, random_seed=random_seed), samples


def compute_obs_obs_similarity(data: np.ndarray, *, method: str = "abs_pearson", reproducible: bool = True) -> np.ndarray:
    """tools/similarity.py:178-246 for a dense cells x genes matrix (no ``top`` / ``bottom``)."""
    if method not in ("pearson", "abs_pearson"):
        raise NotImplementedError("only pearson / abs_pearson are on the hot path (pipeline/direct.py:278)")
    similarity = ut.corrcoef(data, per="row", reproducible=reproducible)
    if method == "abs_pearson":
        np.absolute(similarity, out=similarity)  # similarity.py:214-215
    return similarity


def row_block_of_rank(rows_count: int, rank: int, world_size: int) -> Tuple[int, int]:
    """``[first, last)`` rows of the similarity matrix that ``rank`` computes: contiguous blocks whose sizes
    differ by at most one (the first ``rows_count % world_size`` ranks take the extra row)."""
    assert 0 <= rank < world_size
    base, extra = divmod(rows_count, world_size)
    first = rank * base + min(rank, extra)
    return first, first + base + (1 if rank < extra else 0)


def similarity_top_block(dense_device, first_row: int, last_row: int, top: int, *, method: str = "abs_pearson", stream=None):
    """Rows ``[first_row, last_row)`` of ``top_per(similarity, top)`` on the device: the block of the similarity
    matrix (tensor cores), then per row the ``top`` largest values with ascending column indices
    (top_per.cpp:5-64).  ``dense_device`` is a CUDA float32 ``torch`` tensor [cells, genes] holding ALL rows.
    Returns CUDA tensors ``(indices int32 [rows, top], values float32 [rows, top])``."""
    import torch

    if method not in ("pearson", "abs_pearson"):
        raise NotImplementedError("only pearson / abs_pearson are on the hot path")
    cells, genes = dense_device.shape
    assert dense_device.is_cuda and dense_device.dtype == torch.float32 and dense_device.is_contiguous()
    assert 0 < top < cells, "top must be in 1..cells-1 (top_per.cpp:81-82)"
    rows = last_row - first_row
    handle = ctypes.c_void_p(stream.cuda_stream if stream is not None else torch.cuda.current_stream().cuda_stream)
    block = torch.empty((rows, cells), dtype=torch.float32, device=dense_device.device)
    indices = torch.empty((rows, top), dtype=torch.int32, device=dense_device.device)
    values = torch.empty((rows, top), dtype=torch.float32, device=dense_device.device)
    if rows == 0:
        return indices, values
    library = lib()
    library.mcb_dev_similarity_block_f32.restype = ctypes.c_int
    library.mcb_dev_collect_top_f32.restype = ctypes.c_int
    check(library.mcb_dev_similarity_block_f32(
        ctypes.c_void_p(dense_device.data_ptr() + first_row * genes * 4), ctypes.c_int64(rows), ctypes.c_int64(first_row),
        ctypes.c_void_p(dense_device.data_ptr()), ctypes.c_int64(cells), ctypes.c_int64(genes), ctypes.c_void_p(block.data_ptr()),
        ctypes.c_int64(cells), ctypes.c_int(int(method == "abs_pearson")), handle))
    check(library.mcb_dev_collect_top_f32(
        ctypes.c_int64(top), ctypes.c_void_p(block.data_ptr()), ctypes.c_int64(rows), ctypes.c_int64(cells), ctypes.c_int64(cells),
        ctypes.c_void_p(indices.data_ptr()), ctypes.c_void_p(values.data_ptr()), ctypes.c_int(0), handle))
    return indices, values


def compute_obs_obs_similarity_top(data: np.ndarray, top: int, *, method: str = "abs_pearson", rank: Optional[int] = None,
                                   world_size: Optional[int] = None) -> sp.csr_matrix:
    """``compute_obs_obs_similarity(..., top=top)`` (tools/similarity.py:222-238 -> ``ut.top_per``,
    computation.py:1299-1352) for a pile too large for one GPU's N x N matrices, sharded by row blocks over the
    ranks of the default ``torch.distributed`` group: every rank holds the whole [cells, genes]
    operand, computes its block of rows of the similarity on the tensor cores, keeps the ``top`` largest values of
    each row, and only those (cells x top indices and values) cross NVLink in one NCCL all-gather.  Every rank
    returns the full CSR matrix (float32 data, sorted int32 indices, ``top`` entries per row)."""
    import torch
    import torch.distributed as dist

    if world_size is None:
        world_size = dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1
    if rank is None:
        rank = dist.get_rank() if world_size > 1 else 0
    dense = np.ascontiguousarray(data, dtype="float32")
    cells = dense.shape[0]
    device = torch.device("cuda", torch.cuda.current_device())
    dense_device = torch.from_numpy(dense).to(device)
    first_row, last_row = row_block_of_rank(cells, rank, world_size)
    indices, values = similarity_top_block(dense_device, first_row, last_row, top, method=method)
    if world_size > 1:
        widest = row_block_of_rank(cells, 0, world_size)[1]  # rank 0 owns the (or a) widest block
        padded_indices = torch.zeros((widest, top), dtype=torch.int32, device=device)
        padded_values = torch.zeros((widest, top), dtype=torch.float32, device=device)
        padded_indices[: last_row - first_row] = indices
        padded_values[: last_row - first_row] = values
        all_indices = torch.empty((world_size, widest, top), dtype=torch.int32, device=device)
        all_values = torch.empty((world_size, widest, top), dtype=torch.float32, device=device)
        dist.all_gather_into_tensor(all_indices, padded_indices)
        dist.all_gather_into_tensor(all_values, padded_values)
        sizes = [row_block_of_rank(cells, other, world_size) for other in range(world_size)]
        indices = torch.cat([all_indices[other, : end - start] for other, (start, end) in enumerate(sizes)])
        values = torch.cat([all_values[other, : end - start] for other, (start, end) in enumerate(sizes)])
    indptr = np.arange(cells + 1, dtype="int64") * top  # computation.py:1324
    matrix = sp.csr_matrix((values.cpu().numpy().reshape(-1), indices.cpu().numpy().reshape(-1), indptr), shape=(cells, cells))
    matrix.has_sorted_indices = True
    matrix.has_canonical_format = True
    return matrix


class KnnGraph:
    """Owner of a device-resident ``mcb_graph`` (include/mcb200.h)."""

    def __init__(self, handle: ctypes.c_void_p):
        self._handle = handle

    def __del__(self):
        self.close()

    def close(self) -> None:
        if getattr(self, "_handle", None):
            lib().mcb_graph_free(self._handle)
            self._handle = None

    @property
    def handle(self) -> ctypes.c_void_p:
        return self._handle

    @property
    def k(self) -> int:
        function = lib().mcb_graph_k
        function.restype = ctypes.c_int64
        return int(function(self._handle))

    @property
    def shape(self) -> Tuple[int, int]:
        function = lib().mcb_graph_size
        function.restype = ctypes.c_int64
        size = int(function(self._handle))
        return size, size

    @property
    def nnz(self) -> int:
        function = lib().mcb_graph_nnz
        function.restype = ctypes.c_int64
        return int(function(self._handle))

    def read_start(self):
        """Starts the copy of the graph to freshly allocated host arrays on a stream of the graph's own
        (``mcb_graph_read_start``) and returns them: a worker launches its next pile before ``read_finish``."""
        size, nnz = self.shape[0], self.nnz
        arrays = (np.empty(nnz, dtype="float32"), np.empty(nnz, dtype="float32"), np.empty(nnz, dtype="int32"),
                  np.empty(size + 1, dtype="int32"))
        function = lib().mcb_graph_read_start
        function.restype = ctypes.c_int
        check(function(self._handle, *(ctypes.c_void_p(a.ctypes.data) for a in arrays)))
        return arrays

    def read_finish(self) -> None:
        function = lib().mcb_graph_read_finish
        function.restype = ctypes.c_int
        check(function(self._handle))

    def seed_partition(self, cell_umis: np.ndarray, bounds, max_seeds_count: int, *, random_seed: int,
                       min_seed_size_quantile: float = 0.85, max_seed_size_quantile: float = 0.95,
                       cooldown_pass: float = 0.02, cooldown_node: float = 0.25) -> Tuple[np.ndarray, float, int]:
        """``choose_seeds`` + the first ``optimize_partitions`` of ``compute_candidate_metacells`` on this graph without
        taking it off the device (``mcb_graph_seed_partition``).  ``bounds``: ``(low, target, high)`` UMIs then
        ``(low, target, high)`` sizes.  Returns ``(community_of_nodes, score, seeds_count)``."""
        cell_umis = np.ascontiguousarray(cell_umis, dtype="float32")
        size = self.shape[0]
        if len(cell_umis) != size:
            raise ValueError("cell_umis must have one entry per node")
        communities = np.empty(size, dtype="int32")
        seeds, score = ctypes.c_int64(0), ctypes.c_double(0.0)
        function = lib().mcb_graph_seed_partition
        function.restype = ctypes.c_int
        check(function(self._handle, ctypes.c_void_p(cell_umis.ctypes.data), ctypes.c_uint64(int(random_seed)),
                       ctypes.c_int64(int(max_seeds_count)), ctypes.c_float(min_seed_size_quantile),
                       ctypes.c_float(max_seed_size_quantile), *(ctypes.c_double(float(bound)) for bound in bounds),
                       ctypes.c_double(cooldown_pass), ctypes.c_double(cooldown_node), ctypes.c_double(1 - cooldown_pass),
                       ctypes.c_void_p(communities.ctypes.data), ctypes.byref(seeds), ctypes.byref(score)))
        return communities, score.value, seeds.value

    def to_scipy(self, arrays=None) -> Tuple[sp.csr_matrix, sp.csr_matrix]:
        """``(outgoing_weights, pruned_ranks)`` as CSR float32 / int32 / int32; ``arrays``: what ``read_start`` returned."""
        size = self.shape[0]
        if arrays is None:
            arrays = self.read_start()
        self.read_finish()
        weights, ranks, indices, indptr = arrays
        outgoing_weights = sp.csr_matrix((weights, indices, indptr), shape=(size, size))
        pruned_ranks = sp.csr_matrix((ranks, indices.copy(), indptr.copy()), shape=(size, size))
        for matrix in (outgoing_weights, pruned_ranks):
            matrix.has_sorted_indices = True
            matrix.has_canonical_format = True
        return outgoing_weights, pruned_ranks


def _knn_graph_handle(similarity: np.ndarray, k: int, balanced_ranks_factor: float, incoming_degree_factor: float,
                      outgoing_degree_factor: float, min_outgoing_degree: int) -> KnnGraph:
    if similarity.ndim != 2 or similarity.shape[0] != similarity.shape[1]:
        raise ValueError("similarity must be a square matrix")
    similarity = np.ascontiguousarray(similarity, dtype="float32")
    handle = ctypes.c_void_p()
    function = lib().mcb_host_knn_graph_f32
    function.restype = ctypes.c_int
    check(function(ctypes.c_void_p(host_ptr(similarity, "float32", "similarity")[1]), ctypes.c_int64(similarity.shape[0]),
                   ctypes.c_int64(int(k)), ctypes.c_double(balanced_ranks_factor), ctypes.c_double(incoming_degree_factor),
                   ctypes.c_double(outgoing_degree_factor), ctypes.c_int64(int(min_outgoing_degree)), ctypes.byref(handle)))
    return KnnGraph(handle)


def compute_obs_obs_knn_graph(similarity: np.ndarray, *, k: int,
                              balanced_ranks_factor: float = KNN_BALANCED_RANKS_FACTOR,
                              incoming_degree_factor: float = KNN_INCOMING_DEGREE_FACTOR,
                              outgoing_degree_factor: float = KNN_OUTGOING_DEGREE_FACTOR,
                              min_outgoing_degree: int = KNN_MIN_OUTGOING_DEGREE,
                              return_details: bool = False):
    """tools/knn_graph.py:31-121: the ``obs_outgoing_weights`` CSR matrix of the balanced k-NN graph
    of a symmetric float32 similarity matrix.  With ``return_details`` also the pruned ranks and
    the ``k`` finally used (the reference grows it until every row keeps ``min_outgoing_degree``
    edges, knn_graph.py:251-268)."""
    assert balanced_ranks_factor > 0.0 and incoming_degree_factor > 0.0 and outgoing_degree_factor > 0.0
    assert min_outgoing_degree >= 1
    graph = _knn_graph_handle(similarity, k, balanced_ranks_factor, incoming_degree_factor, outgoing_degree_factor,
                              min_outgoing_degree)
    try:
        outgoing_weights, pruned_ranks = graph.to_scipy()
        used_k = graph.k
    finally:
        graph.close()
    if return_details:
        return outgoing_weights, pruned_ranks, used_k
    return outgoing_weights


def compute_query_atlas_candidates(query_data: np.ndarray, atlas_data: np.ndarray, *,
                                   candidates_count: int = PROJECT_CANDIDATES_COUNT) -> Tuple[np.ndarray, np.ndarray]:
    """The front-end of ``project_query_onto_atlas`` (tools/project.py): the reference computes
    ``query_atlas_corr = ut.cross_corrcoef_rows(query_log_fractions, atlas_log_fractions)`` (:270), a dense
    [query, atlas] matrix in host memory, and then for every query row ``np.argsort(-correlations)`` (:310), of which
    the candidates loop (:317-331) walks a prefix.  Here both steps run on the GPU without the full matrix: for every
    query row the ``candidates_count`` most correlated atlas rows, most correlated first (equal correlations by atlas
    index, as a stable argsort orders them), and their correlations.  Returns ``(indices int32 [query, count],
    correlations float32 [query, count])``; ``indices[:, 0]`` is the reference's ``atlas_anchor_index``.

    The candidates loop skips atlas rows without significant genes (:325-330), so a caller that needs exactly
    ``candidates_count`` accepted candidates asks for a longer prefix."""
    if query_data.ndim != 2 or atlas_data.ndim != 2:
        raise ValueError("query and atlas data must be matrices")
    if query_data.shape[1] != atlas_data.shape[1]:
        raise ValueError("query and atlas data must have the same genes")
    query = np.asarray(query_data, dtype="float32")
    atlas = np.asarray(atlas_data, dtype="float32")
    if query.strides[1] != 4:
        query = np.ascontiguousarray(query)
    if atlas.strides[1] != 4:
        atlas = np.ascontiguousarray(atlas)
    count = min(int(candidates_count), atlas.shape[0])
    if count <= 0:
        raise ValueError("candidates_count must be positive and the atlas must not be empty")
    indices = np.empty((query.shape[0], count), dtype="int32")
    correlations = np.empty((query.shape[0], count), dtype="float32")
    call("mcb_host_project_top_f32", host_ptr(query, "float32", "query"), ("i64", query.shape[0]), ("i64", row_stride(query)),
         host_ptr(atlas, "float32", "atlas"), ("i64", atlas.shape[0]), ("i64", row_stride(atlas)), ("i64", query.shape[1]),
         ("i64", count), host_ptr(indices, "int32", "indices", True), host_ptr(correlations, "float32", "correlations", True))
    return indices, correlations


def _both_layouts(edge_weights):
    """The outgoing CSR and the incoming CSC of a graph (``ut.to_layout(..., "column_major")`` of candidates.py:696)."""
    from . import extensions as xt  # noqa: F401  (fail early when the library is missing)

    outgoing = sp.csr_matrix(edge_weights)
    if outgoing.dtype != np.float32 or outgoing.indices.dtype != np.int32 or outgoing.indptr.dtype != np.int32:
        outgoing = sp.csr_matrix((outgoing.data.astype("float32"), outgoing.indices.astype("int32"),
                                  outgoing.indptr.astype("int32")), shape=outgoing.shape)
    if outgoing.shape[0] != outgoing.shape[1]:
        raise ValueError("edge_weights must be square")
    outgoing.sort_indices()
    incoming = ut.relayout_compressed(outgoing)  # the CSR of the transpose = the CSC of the graph, on the GPU
    return outgoing, incoming


def choose_seeds(*, edge_weights, seed_of_cells: Optional[np.ndarray] = None, max_seeds_count: int,
                 min_seed_size_quantile: float = MIN_SEED_SIZE_QUANTILE, max_seed_size_quantile: float = MAX_SEED_SIZE_QUANTILE,
                 random_seed: int) -> np.ndarray:
    """tools/candidates.py:653-716: the initial assignment of cells to seeds.  ``seed_of_cells`` (int32, negative =
    no seed), when given, is completed in place and returned; otherwise a new vector is."""
    from . import extensions as xt

    outgoing, incoming = _both_layouts(edge_weights)
    size = outgoing.shape[0]
    if seed_of_cells is None:
        seed_of_cells = np.full(size, -1, dtype="int32")
    elif seed_of_cells.dtype != np.int32 or len(seed_of_cells) != size:
        raise ValueError("seed_of_cells must be int32 with one entry per cell")
    xt.choose_seeds(outgoing.data, outgoing.indices, outgoing.indptr, incoming.data, incoming.indices, incoming.indptr,
                    random_seed, max_seeds_count, min_seed_size_quantile, max_seed_size_quantile, seed_of_cells)
    return seed_of_cells


def optimize_partitions(*, edge_weights, community_of_nodes: np.ndarray, node_umis: np.ndarray, low_partition_umis: int,
                        target_partition_umis: int, high_partition_umis: int, low_partition_size: int,
                        target_partition_size: int, high_partition_size: int, cooldown_pass: float = COOLDOWN_PASS,
                        cooldown_node: float = COOLDOWN_NODE, random_seed: int) -> float:
    """tools/candidates.py:752-830: ``community_of_nodes`` is optimised in place, the score is returned (every
    community hot, ``cold_temperature = cooldown_pass`` as :823-829)."""
    from . import extensions as xt

    if not (0 < low_partition_size < target_partition_size < high_partition_size):
        raise ValueError("partition sizes must be 0 < low < target < high")
    if not (0 < low_partition_umis < target_partition_umis < high_partition_umis):
        raise ValueError("partition UMIs must be 0 < low < target < high")
    outgoing, incoming = _both_layouts(edge_weights)
    communities_count = int(np.max(community_of_nodes)) + 1
    if communities_count <= 0:
        raise ValueError("community_of_nodes holds no community")
    hot_mask = np.ones(max(communities_count, 2), dtype="int8")
    return xt.optimize_partitions(outgoing.data, outgoing.indices, outgoing.indptr, incoming.data, incoming.indices,
                                  incoming.indptr, random_seed, np.ascontiguousarray(node_umis, dtype="float32"),
                                  float(low_partition_umis), float(target_partition_umis), float(high_partition_umis),
                                  float(low_partition_size), float(target_partition_size), float(high_partition_size),
                                  cooldown_pass, cooldown_node, community_of_nodes, hot_mask, cooldown_pass)
