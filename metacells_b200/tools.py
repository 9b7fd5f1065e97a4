"""Host-side mirror of the hot-path entry points of ``metacells.tools`` on numpy / scipy data
(AnnData is not needed: ``what`` is the matrix itself and results are returned instead of stored):

* ``downsample_cells``            tools/downsample.py:25-118
* ``compute_obs_obs_similarity``  tools/similarity.py:29-117, 178-246 (``abs_pearson`` / ``pearson``)
* ``compute_obs_obs_knn_graph``   tools/knn_graph.py:31-121, 203-283

Defaults are the reference's (metacells/parameters.py).
"""
from __future__ import annotations

import ctypes
from typing import Optional, Tuple

import numpy as np
import scipy.sparse as sp

from . import utilities as ut
from ._lib import check, host_ptr, lib

__all__ = ["downsample_cells", "compute_obs_obs_similarity", "compute_obs_obs_knn_graph", "KnnGraph"]

# metacells/parameters.py
DOWNSAMPLE_MIN_SAMPLES = 750
DOWNSAMPLE_MIN_CELL_QUANTILE = 0.05
DOWNSAMPLE_MAX_CELL_QUANTILE = 0.5
KNN_BALANCED_RANKS_FACTOR = float(np.sqrt(10))
KNN_INCOMING_DEGREE_FACTOR = 3.0
KNN_OUTGOING_DEGREE_FACTOR = 1.0
KNN_MIN_OUTGOING_DEGREE = 2


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

    def to_scipy(self) -> Tuple[sp.csr_matrix, sp.csr_matrix]:
        """``(outgoing_weights, pruned_ranks)`` as CSR float32 / int32 / int32."""
        size, nnz = self.shape[0], self.nnz
        weights = np.empty(nnz, dtype="float32")
        ranks = np.empty(nnz, dtype="float32")
        indices = np.empty(nnz, dtype="int32")
        indptr = np.empty(size + 1, dtype="int32")
        function = lib().mcb_graph_read
        function.restype = ctypes.c_int
        check(function(self._handle, *(ctypes.c_void_p(a.ctypes.data) for a in (weights, ranks, indices, indptr))))
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
