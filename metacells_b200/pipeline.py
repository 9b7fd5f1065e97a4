"""The per-pile k-NN graph of ``compute_direct_metacells`` (pipeline/direct.py:26-283), device
resident: the UMI matrix is uploaded once, downsampled, sliced to the feature genes, log-transformed,
correlated, ranked, balanced and pruned in HBM, and only the graph comes back.

``Pile`` wraps the ``mcb_pile`` handle of include/mcb200.h.  Feature-gene selection
(pipeline/select.py) is outside the hot path: the caller passes the gene indices, or uses the
``top_total_genes`` stand-in (the genes with the largest downsampled totals).
"""
from __future__ import annotations

import ctypes
from typing import Iterable, Iterator, Optional, Tuple

import numpy as np
import scipy.sparse as sp

from . import tools as tl
from ._lib import check, lib

__all__ = ["Pile", "compute_pile_knn_graph", "knn_k_of_pile", "log_fraction_table", "compute_pile_seed_partition",
           "compute_piles_seed_partitions", "compute_piles_candidates", "metacell_bounds", "seeds_count_for"]

# metacells/parameters.py
CELLS_SIMILARITY_VALUE_REGULARIZATION = 1 / 7
TARGET_METACELL_SIZE = 48
TARGET_METACELL_UMIS = 160000
KNN_K_SIZE_FACTOR = 2.0
KNN_K_UMIS_QUANTILE = 0.1
MIN_KNN_K = 30
MIN_METACELL_SIZE = 12
CANDIDATES_MIN_SPLIT_SIZE_FACTOR = 2.0
CANDIDATES_MAX_MERGE_SIZE_FACTOR = 0.5


def log_fraction_table(max_count: int, regularization: float = CELLS_SIMILARITY_VALUE_REGULARIZATION) -> np.ndarray:
    """``np.log2(float32(c) + regularization)`` for c = 0 .. max_count, computed by numpy itself in
    float32 exactly as direct.py:245-249 does element by element."""
    table = np.arange(max_count + 1, dtype="float32")
    table += regularization
    np.log2(table, out=table)
    return table


def knn_k_of_pile(cells_count: int, cell_umis: np.ndarray, *, target_metacell_size: int = TARGET_METACELL_SIZE,
                  target_metacell_umis: int = TARGET_METACELL_UMIS, knn_k_size_factor: float = KNN_K_SIZE_FACTOR,
                  knn_k_umis_quantile: float = KNN_K_UMIS_QUANTILE, min_knn_k: Optional[int] = MIN_KNN_K) -> int:
    """pipeline/direct.py:251-267."""
    by_size = int(round(cells_count / target_metacell_size))
    by_median = int(round(knn_k_size_factor * target_metacell_umis / float(np.median(cell_umis))))
    by_quantile = int(round(target_metacell_umis / np.quantile(cell_umis, knn_k_umis_quantile)))
    return max(by_size, by_median, by_quantile, min_knn_k or 0)


def _ptr(array: np.ndarray) -> ctypes.c_void_p:
    return ctypes.c_void_p(array.ctypes.data)


_POOLS: dict = {}


def worker_pool(workers: int):
    """The host threads that drive piles side by side on the calling thread's device, kept for the life of the process:
    every thread owns a CUDA stream, a page-locked staging slot and a cached log table in the library (runtime.cu), so
    a fresh executor per call would leak them.  One pool per (device, size); ``map`` may be used from one caller at a
    time."""
    from concurrent.futures import ThreadPoolExecutor

    device = ctypes.c_int(0)
    check(lib().mcb_current_device(ctypes.byref(device)))
    key = (device.value, int(workers))
    pool = _POOLS.get(key)
    if pool is None:
        def on_device(index=device.value) -> None:  # the current device is per thread
            check(lib().mcb_set_device(ctypes.c_int(index)))

        pool = _POOLS[key] = ThreadPoolExecutor(max_workers=int(workers), initializer=on_device,
                                                thread_name_prefix=f"mcb-pile-{device.value}")
    return pool


def _side_by_side(workers: Optional[int]) -> int:
    """How many one-SM optimisations run at once: one per SM unless the caller says otherwise."""
    if workers is not None:
        return max(1, int(workers))
    count = ctypes.c_int(0)
    check(lib().mcb_device_sm_count(ctypes.byref(count)))
    return count.value


class CompactPile:
    """The compact host copy of a pile (``compact_csr``): ``indices`` uint16, ``counts`` uint8 (255: see the escapes),
    ``indptr`` int32, ``escape_at`` int64 (ascending positions) and ``escape_values`` float32, ``shape``."""

    def __init__(self, indices, counts, indptr, escape_at, escape_values, shape):
        self.indices, self.counts, self.indptr = indices, counts, indptr
        self.escape_at, self.escape_values, self.shape = escape_at, escape_values, shape

    @property
    def nbytes(self) -> int:
        return sum(array.nbytes for array in (self.indices, self.counts, self.indptr, self.escape_at, self.escape_values))


def compact_csr(umis: sp.csr_matrix) -> CompactPile:
    """A UMI matrix (CSR, at most 65 536 genes) in the 3-bytes-per-non-zero form ``Pile.from_compact`` uploads: 16-bit gene
    indices, 8-bit counts; counts of 255 and more, and counts that are not whole numbers, go to a list of escapes."""
    umis = sp.csr_matrix(umis)
    if umis.shape[1] > 65536:
        raise ValueError("a compact pile has at most 65536 genes")
    data = np.asarray(umis.data, dtype="float32")
    escaped = ~((data >= 0) & (data < 255) & (data == np.floor(data)))
    counts = np.where(escaped, 255, data).astype("uint8")
    escape_at = np.flatnonzero(escaped).astype("int64")
    return CompactPile(np.ascontiguousarray(umis.indices.astype("uint16")), np.ascontiguousarray(counts),
                       np.ascontiguousarray(umis.indptr, dtype="int32"), escape_at,
                       np.ascontiguousarray(data[escape_at], dtype="float32"), umis.shape)


class Pile:
    """A device-resident pile (cells x genes UMI CSR matrix)."""

    def __init__(self, umis: sp.csr_matrix, *, max_cell_total: Optional[float] = None, pin: bool = False):
        """``max_cell_total``: the largest total of a cell when the caller has the totals anyway
        (tools/downsample.py:84-90 needs them for the samples count); lets the downsampling start chunk by chunk
        under the upload without a pass over the data.  ``pin``: page-lock the arrays for the life of the pile, so that
        the constructor returns while the upload is still running (see ``compute_piles_knn_graphs``)."""
        umis = sp.csr_matrix(umis)
        self._data = np.ascontiguousarray(umis.data, dtype="float32")
        self._indices = np.ascontiguousarray(umis.indices, dtype="int32")
        self._indptr = np.ascontiguousarray(umis.indptr, dtype="int32")
        self.shape = umis.shape
        self._pinned = []
        if pin:
            for array in (self._data, self._indices, self._indptr):
                if array.nbytes:
                    check(lib().mcb_host_pin(_ptr(array), ctypes.c_uint64(array.nbytes)))
                    self._pinned.append(array)
        self._handle = ctypes.c_void_p()
        function = lib().mcb_pile_create_host
        function.restype = ctypes.c_int
        check(function(_ptr(self._data), _ptr(self._indices), _ptr(self._indptr), ctypes.c_int64(umis.shape[0]),
                       ctypes.c_int64(umis.shape[1]), ctypes.byref(self._handle)))
        self.samples: Optional[int] = None
        self.gene_top_threshold: Optional[float] = None
        if max_cell_total is not None:
            function = lib().mcb_pile_set_max_cell_total
            function.restype = ctypes.c_int
            check(function(self._handle, ctypes.c_uint64(int(max_cell_total))))

    @classmethod
    def _adopt(cls, handle: ctypes.c_void_p, shape: Tuple[int, int]) -> "Pile":
        """A pile whose arrays live on the device only (gathered from an atlas, or made from a dense device matrix)."""
        pile = cls.__new__(cls)
        pile._data = pile._indices = pile._indptr = None
        pile.shape = shape
        pile._pinned = []
        pile._handle = handle
        pile.samples = None
        pile.gene_top_threshold = None
        return pile

    @classmethod
    def from_compact(cls, compact: "CompactPile", *, max_cell_total: Optional[float] = None) -> "Pile":
        """A pile uploaded from its compact host copy (``compact_csr``): 3 bytes per non-zero on the host link instead of
        8 (``mcb_pile_create_host_compact``); on the device it is the pile ``Pile(umis)`` makes."""
        pile = cls._adopt(ctypes.c_void_p(), compact.shape)
        pile._compact = compact  # the arrays stay alive while the asynchronous upload runs
        function = lib().mcb_pile_create_host_compact
        function.restype = ctypes.c_int
        check(function(_ptr(compact.indices), _ptr(compact.counts), _ptr(compact.indptr), ctypes.c_int64(compact.shape[0]),
                       ctypes.c_int64(compact.shape[1]), _ptr(compact.escape_at), _ptr(compact.escape_values),
                       ctypes.c_int64(len(compact.escape_at)), ctypes.byref(pile._handle)))
        if max_cell_total is not None:
            pile.set_max_cell_total(max_cell_total)
        return pile

    @classmethod
    def from_atlas(cls, atlas_handle: ctypes.c_void_p, rows: np.ndarray, genes: int,
                   max_cell_total: Optional[float] = None) -> "Pile":
        """The pile of the atlas rows ``rows`` (``ut.slice(adata, obs=...)`` of divide_and_conquer.py:1267), gathered on
        the device (``mcb_pile_create_from_atlas``): only the row list crosses the host link."""
        rows = np.ascontiguousarray(rows, dtype="int64")
        handle = ctypes.c_void_p()
        function = lib().mcb_pile_create_from_atlas
        function.restype = ctypes.c_int
        check(function(atlas_handle, _ptr(rows), ctypes.c_int64(len(rows)), ctypes.byref(handle)))
        pile = cls._adopt(handle, (len(rows), int(genes)))
        if max_cell_total is not None:
            pile.set_max_cell_total(max_cell_total)
        return pile

    @classmethod
    def from_dense_device(cls, pointer: int, cells: int, genes: int) -> "Pile":
        """The pile of a dense float32 [cells, genes] device matrix (``mcb_pile_create_dense_dev``)."""
        handle = ctypes.c_void_p()
        function = lib().mcb_pile_create_dense_dev
        function.restype = ctypes.c_int
        check(function(ctypes.c_void_p(pointer), ctypes.c_int64(cells), ctypes.c_int64(genes), ctypes.byref(handle)))
        return cls._adopt(handle, (int(cells), int(genes)))

    def set_max_cell_total(self, max_cell_total: float) -> None:
        """``mcb_pile_set_max_cell_total``: the largest total of a cell, when the caller has the totals."""
        function = lib().mcb_pile_set_max_cell_total
        function.restype = ctypes.c_int
        check(function(self._handle, ctypes.c_uint64(int(max_cell_total))))

    def _largest_count(self) -> int:
        """An upper bound of the values the log table must cover: a downsampled cell holds at most ``samples`` UMIs."""
        if self.samples is not None:
            return int(self.samples)
        if self._data is None:
            raise ValueError("max_count is needed for a device pile that was not downsampled")
        return int(self._data.max(initial=0))

    def cell_totals(self) -> np.ndarray:
        """The total UMIs of every cell (float32, ``umis.sum(axis=1)``), for a pile made from host arrays; a pile gathered
        from an atlas takes them from ``Atlas.cell_umis``."""
        if self._data is None:
            raise ValueError("the pile has no host copy of its UMIs: pass the totals (Atlas.cell_umis()[rows])")
        matrix = sp.csr_matrix((self._data, self._indices, self._indptr), shape=self.shape)
        return np.asarray(matrix.sum(axis=1)).reshape(-1).astype("float32", copy=False)

    def wait_upload(self) -> None:
        """Returns when the pile's upload has arrived (``mcb_pile_wait_upload``)."""
        check(lib().mcb_pile_wait_upload(self._handle))

    def close(self) -> None:
        if getattr(self, "_handle", None):
            lib().mcb_pile_free(self._handle)  # waits for the upload
            self._handle = None
        for array in getattr(self, "_pinned", []):
            lib().mcb_host_unpin(_ptr(array))
        self._pinned = []

    def __del__(self):
        self.close()

    def __enter__(self):
        return self

    def __exit__(self, *_):
        self.close()

    def downsample(self, samples: int, random_seed: int) -> None:
        """tools/downsample.py:99."""
        function = lib().mcb_pile_downsample
        function.restype = ctypes.c_int
        check(function(self._handle, ctypes.c_int32(int(samples)), ctypes.c_uint64(int(random_seed))))
        self.samples = int(samples)

    def downsampled(self) -> sp.csr_matrix:
        compact = getattr(self, "_compact", None)
        if self._data is None and compact is None:
            raise NotImplementedError("a pile gathered on the device keeps no host copy of its structure")
        indices = self._indices.copy() if compact is None else compact.indices.astype("int32")
        indptr = (self._indptr if compact is None else compact.indptr).copy()
        data = np.empty(len(indices), dtype="float32")
        function = lib().mcb_pile_read_downsampled
        function.restype = ctypes.c_int
        check(function(self._handle, _ptr(data)))
        return sp.csr_matrix((data, indices, indptr), shape=self.shape)

    def want_gene_statistics(self, min_top_value: float) -> None:
        """Before ``downsample``: have its kernel add up the per-gene statistics the feature-gene selection starts from
        (``mcb_pile_want_gene_statistics``); ``min_top_value`` is ``min_gene_top3`` of pipeline/select.py."""
        function = lib().mcb_pile_want_gene_statistics
        function.restype = ctypes.c_int
        check(function(self._handle, ctypes.c_float(float(min_top_value))))
        self.gene_top_threshold = float(min_top_value)

    def gene_statistics(self) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
        """``(totals float32, sums of squares uint64, cells at or above the top value int32)`` per gene of the downsampled
        pile (``mcb_pile_gene_statistics``)."""
        genes = self.shape[1]
        totals, squares, top = np.empty(genes, dtype="float32"), np.empty(genes, dtype="uint64"), np.empty(genes, dtype="int32")
        function = lib().mcb_pile_gene_statistics
        function.restype = ctypes.c_int
        check(function(self._handle, _ptr(totals), _ptr(squares), _ptr(top)))
        return totals, squares, top

    def gene_columns(self, genes: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
        """The non-zero values of ``genes`` in cell order, ``(values float32 [len(genes), cells], sizes int32)``
        (``mcb_pile_gene_columns``): row g holds ``sizes[g]`` values."""
        genes = np.ascontiguousarray(genes, dtype="int32")
        values = np.zeros((len(genes), self.shape[0]), dtype="float32")
        sizes = np.zeros(len(genes), dtype="int32")
        function = lib().mcb_pile_gene_columns
        function.restype = ctypes.c_int
        check(function(self._handle, _ptr(genes), ctypes.c_int64(len(genes)), _ptr(values), _ptr(sizes)))
        return values, sizes

    def gene_totals(self) -> np.ndarray:
        totals = np.empty(self.shape[1], dtype="float32")
        function = lib().mcb_pile_gene_totals
        function.restype = ctypes.c_int
        check(function(self._handle, _ptr(totals)))
        return totals

    def top_total_genes(self, features: int) -> np.ndarray:
        """The ``features`` genes with the largest totals (ties: lower gene index), ascending."""
        totals = self.gene_totals()
        return np.sort(np.argsort(-totals, kind="stable")[:features]).astype("int32")

    def select_top_genes(self, features: int) -> np.ndarray:
        """As ``top_total_genes``, selected on the device (``mcb_pile_top_genes``); the list also stays there, so
        ``knn_graph(None, ...)`` uses it without another upload."""
        genes = np.empty(features, dtype="int32")
        function = lib().mcb_pile_top_genes
        function.restype = ctypes.c_int
        check(function(self._handle, ctypes.c_int64(features), _ptr(genes)))
        return genes

    def log_fraction(self, feature_genes: Optional[np.ndarray], *,
                     regularization: float = CELLS_SIMILARITY_VALUE_REGULARIZATION,
                     max_count: Optional[int] = None) -> np.ndarray:
        """direct.py:242-249: the dense float32 [cells, features] operand ``log2(downsampled[:, genes] + 1/7)``
        (``mcb_pile_log_fraction``); ``feature_genes=None`` uses the list ``select_top_genes`` left on the device."""
        if feature_genes is not None:
            feature_genes = np.ascontiguousarray(feature_genes, dtype="int32")
            features = len(feature_genes)
        else:
            features = 0
        if max_count is None:
            max_count = self._largest_count()
        table = log_fraction_table(min(max_count, 1 << 20), regularization)
        if feature_genes is None:
            function = lib().mcb_pile_top_genes_count
            function.restype = ctypes.c_int64
            features = int(function(self._handle))
        dense = np.empty((self.shape[0], features), dtype="float32")
        function = lib().mcb_pile_log_fraction
        function.restype = ctypes.c_int
        check(function(self._handle, _ptr(feature_genes) if feature_genes is not None else None, ctypes.c_int64(features),
                       _ptr(table), ctypes.c_int64(len(table)), ctypes.c_float(np.float32(regularization)), _ptr(dense), None))
        return dense

    def knn_graph(self, feature_genes: Optional[np.ndarray], *, k: int, method: str = "abs_pearson",
                  regularization: float = CELLS_SIMILARITY_VALUE_REGULARIZATION,
                  balanced_ranks_factor: float = tl.KNN_BALANCED_RANKS_FACTOR,
                  incoming_degree_factor: float = tl.KNN_INCOMING_DEGREE_FACTOR,
                  outgoing_degree_factor: float = tl.KNN_OUTGOING_DEGREE_FACTOR,
                  min_outgoing_degree: int = tl.KNN_MIN_OUTGOING_DEGREE, max_count: Optional[int] = None) -> tl.KnnGraph:
        """direct.py:242-283: log-fraction of the feature genes, similarity, balanced k-NN graph."""
        if method not in ("pearson", "abs_pearson"):
            raise NotImplementedError("only pearson / abs_pearson are on the hot path")
        if feature_genes is not None:
            feature_genes = np.ascontiguousarray(feature_genes, dtype="int32")
        if max_count is None:
            max_count = self._largest_count()
        table = log_fraction_table(min(max_count, 1 << 20), regularization)
        handle = ctypes.c_void_p()
        outside = ctypes.c_int(0)
        function = lib().mcb_pile_knn_graph
        function.restype = ctypes.c_int
        check(function(self._handle, _ptr(feature_genes) if feature_genes is not None else None,
                       ctypes.c_int64(len(feature_genes) if feature_genes is not None else 0), _ptr(table),
                       ctypes.c_int64(len(table)), ctypes.c_float(np.float32(regularization)),
                       ctypes.c_int(int(method == "abs_pearson")), ctypes.c_int64(int(k)), ctypes.c_double(balanced_ranks_factor),
                       ctypes.c_double(incoming_degree_factor), ctypes.c_double(outgoing_degree_factor),
                       ctypes.c_int64(int(min_outgoing_degree)), ctypes.byref(outside), ctypes.byref(handle)))
        graph = tl.KnnGraph(handle)
        graph.outside_table = bool(outside.value)
        return graph


def _start_selection(pile: "Pile", features) -> None:
    """Before ``pile.downsample``: the reference's gene selection needs the per-gene statistics the kernel adds up."""
    if features == "select":
        from . import select

        pile.want_gene_statistics(select.SELECT_MIN_GENE_TOP3)


def _feature_genes_of(pile: "Pile", features) -> Optional[np.ndarray]:
    """The feature genes of a downsampled pile.  ``features == "select"``: ``extract_selected_data`` of pipeline/select.py
    (``select.select_feature_genes``: high top-3, high total, high relative variance), returned as gene indices for
    ``knn_graph``; an integer: the stand-in of SURVEY.md 8d, that many genes of the largest totals, selected on the device
    (the list stays there: ``knn_graph(None, ...)``)."""
    if features == "select":
        from . import select

        return np.flatnonzero(select.select_feature_genes(pile)).astype("int32")
    pile.select_top_genes(min(int(features), pile.shape[1] - 1))
    return None


def compute_pile_knn_graph(umis: sp.csr_matrix, *, features=2000, feature_genes: Optional[np.ndarray] = None,
                           k: Optional[int] = None, random_seed: int = 123456, upload_turn=None) -> Tuple[sp.csr_matrix, dict]:
    """The k-NN graph part of ``compute_direct_metacells`` for one pile: downsample (tools/downsample.py),
    feature genes, log-fraction, ``abs_pearson`` similarity, ``compute_obs_obs_knn_graph``.
    Returns ``(obs_outgoing_weights, details)``."""
    cell_umis = np.asarray(umis.sum(axis=1)).reshape(-1).astype("float32")  # direct.py:214-218
    samples = int(round(min(max(tl.DOWNSAMPLE_MIN_SAMPLES, np.quantile(cell_umis, tl.DOWNSAMPLE_MIN_CELL_QUANTILE)),
                            np.quantile(cell_umis, tl.DOWNSAMPLE_MAX_CELL_QUANTILE))))
    if k is None:
        k = knn_k_of_pile(umis.shape[0], cell_umis)
    if upload_turn is None:  # a lock shared by the threads that walk piles side by side: one upload at a time
        pile = Pile(umis, max_cell_total=float(cell_umis.max()))
    else:
        with upload_turn:
            pile = Pile(umis, max_cell_total=float(cell_umis.max()))
            pile.wait_upload()
    with pile:
        if feature_genes is None:
            _start_selection(pile, features)
        pile.downsample(samples, random_seed)
        selected_on_device = feature_genes is None and features != "select"
        if feature_genes is None:
            feature_genes = _feature_genes_of(pile, features)
            if feature_genes is None:
                feature_genes = pile.select_top_genes(features)
        graph = pile.knn_graph(None if selected_on_device else feature_genes, k=k)
        try:
            weights, pruned = graph.to_scipy()
            details = {"k": graph.k, "samples": samples, "feature_genes": feature_genes, "pruned_ranks": pruned}
        finally:
            graph.close()
    return weights, details


def compute_piles_knn_graphs(piles: Iterable[sp.csr_matrix], *, features: int = 2000, random_seed: int = 123456,
                             pin: bool = False, workers: int = 1) -> Iterator[Tuple[sp.csr_matrix, dict]]:
    """``compute_pile_knn_graph`` over the piles of one GPU, one after the other as ``ut.parallel_map`` hands them to a
    worker (pipeline/divide_and_conquer.py:1264-1295; every pile gets the same ``random_seed``, :1274-1279), with the
    upload of pile k + 1 started before the graph of pile k is asked for: it runs on the pile's copy stream under the
    kernels of pile k when the pile's arrays are page-locked (``mcb_pile_create_host`` copies from pageable memory
    synchronously).  ``pin=True`` page-locks the arrays of every pile for its lifetime; that costs more than it saves
    unless the arrays are reused, so a caller that streams piles keeps them in page-locked buffers of its own.
    Yields ``(obs_outgoing_weights, details)`` per pile, in order; results do not depend on the overlap.

    ``workers > 1``: that many host threads walk the piles instead (``compute_pile_knn_graph`` each, on the thread's own
    stream), so one pile's upload, read-back and host-side steps lie under another pile's kernels; at most
    ``workers + 1`` piles are taken from ``piles`` ahead of the one being yielded."""
    if workers > 1:
        from collections import deque
        import threading

        upload_turn = threading.Lock()  # the threads take turns uploading: each upload under the others' kernels
        pool = worker_pool(int(workers))
        in_flight: deque = deque()
        for umis in piles:
            in_flight.append(pool.submit(compute_pile_knn_graph, umis, features=features, random_seed=random_seed,
                                         upload_turn=upload_turn))
            if len(in_flight) > workers:
                yield in_flight.popleft().result()
        while in_flight:
            yield in_flight.popleft().result()
        return

    def start(umis: sp.csr_matrix):
        cell_umis = np.asarray(umis.sum(axis=1)).reshape(-1).astype("float32")
        return Pile(umis, max_cell_total=float(cell_umis.max(initial=0)), pin=pin), cell_umis

    iterator = iter(piles)
    try:
        pending = start(next(iterator))
    except StopIteration:
        return
    while pending is not None:
        pile, cell_umis = pending
        try:
            following = next(iterator, None)
            pending = start(following) if following is not None else None
            samples = int(round(min(max(tl.DOWNSAMPLE_MIN_SAMPLES, np.quantile(cell_umis, tl.DOWNSAMPLE_MIN_CELL_QUANTILE)),
                                    np.quantile(cell_umis, tl.DOWNSAMPLE_MAX_CELL_QUANTILE))))
            k = knn_k_of_pile(pile.shape[0], cell_umis)
            pile.downsample(samples, random_seed)
            feature_genes = pile.select_top_genes(features)
            graph = pile.knn_graph(None, k=k)
            try:
                weights, pruned = graph.to_scipy()
                details = {"k": graph.k, "samples": samples, "feature_genes": feature_genes, "pruned_ranks": pruned}
            finally:
                graph.close()
        finally:
            pile.close()
        yield weights, details


def metacell_bounds(*, target_metacell_size: int = TARGET_METACELL_SIZE, target_metacell_umis: int = TARGET_METACELL_UMIS,
                    min_metacell_size: int = MIN_METACELL_SIZE, min_split_size_factor: float = CANDIDATES_MIN_SPLIT_SIZE_FACTOR,
                    max_merge_size_factor: float = CANDIDATES_MAX_MERGE_SIZE_FACTOR):
    """tools/candidates.py:151-160: ``(min_metacell_umis, target_metacell_umis, max_metacell_umis, min_metacell_size,
    target_metacell_size, max_metacell_size)``, the bounds ``optimize_partitions`` takes."""
    import math

    min_umis = int(math.floor(target_metacell_umis * max_merge_size_factor))
    max_umis = int(math.ceil(target_metacell_umis * min_split_size_factor)) - 1
    min_size = max(min_metacell_size, int(math.floor(target_metacell_size * max_merge_size_factor)))
    max_size = max(int(math.ceil(target_metacell_size * min_split_size_factor)) - 1, min_size + 1)
    return float(min_umis), float(target_metacell_umis), float(max_umis), float(min_size), float(target_metacell_size), float(max_size)


def seeds_count_for(total_umis: float, total_size: int, bounds) -> int:
    """tools/candidates.py:286-322 (``_seeds_count_for``)."""
    import math

    min_umis, _, max_umis, min_size, _, max_size = bounds
    least_by_umis, most_by_umis = int(math.ceil(total_umis / max_umis)), int(math.ceil(total_umis / min_umis))
    least_by_size, most_by_size = int(math.ceil(total_size / max_size)), int(math.ceil(total_size / min_size))
    if most_by_size < least_by_umis:
        return int(round((most_by_size + least_by_umis) / 2))
    if most_by_umis < least_by_size:
        return int(round((most_by_umis + least_by_size) / 2))
    return int(round((max(least_by_size, least_by_umis) + min(most_by_size, most_by_umis)) / 2))


def compute_pile_seed_partition(edge_weights: sp.csr_matrix, cell_umis: np.ndarray, *, random_seed: int = 123456,
                                min_seed_size_quantile: float = tl.MIN_SEED_SIZE_QUANTILE,
                                max_seed_size_quantile: float = tl.MAX_SEED_SIZE_QUANTILE,
                                cooldown_pass: float = tl.COOLDOWN_PASS, cooldown_node: float = tl.COOLDOWN_NODE,
                                **bounds_arguments) -> Tuple[np.ndarray, float, int]:
    """The opening of ``compute_candidate_metacells`` for one pile (tools/candidates.py:128-196 and the first
    ``_optimize_partitions`` of ``_reduce_communities``, :351-368): the seeds count from the pile's UMIs, ``choose_seeds``,
    then ``optimize_partitions`` with every community hot and ``cold_temperature = 1 - cooldown_pass``.  Both run on the
    GPU, one CTA each, on the calling thread's own stream.  Returns ``(community_of_nodes, score, seeds_count)``.
    The split / min-cut rounds that follow in the reference (:370-407; ``igraph``) stay with the caller."""
    from . import extensions as xt

    outgoing, incoming = tl._both_layouts(edge_weights)
    cell_umis = np.ascontiguousarray(cell_umis, dtype="float32")
    bounds = metacell_bounds(**bounds_arguments)
    total_umis = float(sum(cell_umis))  # candidates.py:162: Python's left-to-right sum of float32 values
    wanted = seeds_count_for(total_umis, outgoing.shape[0], bounds)
    community_of_nodes = np.full(outgoing.shape[0], -1, dtype="int32")
    graph = (outgoing.data, outgoing.indices, outgoing.indptr, incoming.data, incoming.indices, incoming.indptr)
    seeds_count = xt.choose_seeds(*graph, random_seed, wanted, min_seed_size_quantile, max_seed_size_quantile,
                                  community_of_nodes)
    hot_mask = np.ones(max(seeds_count, 2), dtype="int8")
    score = xt.optimize_partitions(*graph, random_seed, cell_umis, *bounds, cooldown_pass, cooldown_node, community_of_nodes,
                                   hot_mask, 1 - cooldown_pass)
    return community_of_nodes, score, seeds_count


def compute_piles_seed_partitions(piles, *, workers: Optional[int] = None, **arguments):
    """``compute_pile_seed_partition`` for ``(edge_weights, cell_umis)`` pairs, side by side: the seed chooser and the
    partition optimiser are one persistent CTA each, so every pile takes one SM and up to ``workers`` (default: one per
    SM, at most one per pile) run at once, each on the stream of the host thread that drives it (the entry points
    release the GIL).  The reference runs its piles in forked processes, one core each
    (pipeline/divide_and_conquer.py:1291-1295).  Results are in the order of ``piles`` and do not depend on ``workers``."""
    piles = list(piles)
    if not piles:
        return []
    pool = worker_pool(_side_by_side(workers))
    return list(pool.map(lambda pile: compute_pile_seed_partition(pile[0], pile[1], **arguments), piles))


def pile_knn_graph_on_device(umis: sp.csr_matrix, *, features=2000, random_seed: int = 123456):
    """``compute_pile_knn_graph`` that leaves the graph on the device: ``(KnnGraph, cell_umis)``; the caller closes the
    graph."""
    cell_umis = np.asarray(umis.sum(axis=1)).reshape(-1).astype("float32")
    samples = int(round(min(max(tl.DOWNSAMPLE_MIN_SAMPLES, np.quantile(cell_umis, tl.DOWNSAMPLE_MIN_CELL_QUANTILE)),
                            np.quantile(cell_umis, tl.DOWNSAMPLE_MAX_CELL_QUANTILE))))
    k = knn_k_of_pile(umis.shape[0], cell_umis)
    with Pile(umis, max_cell_total=float(cell_umis.max(initial=0))) as pile:
        _start_selection(pile, features)
        pile.downsample(samples, random_seed)
        return pile.knn_graph(_feature_genes_of(pile, features), k=k), cell_umis


def compute_piles_candidates(piles: Iterable[sp.csr_matrix], *, features: int = 2000, random_seed: int = 123456,
                             workers: Optional[int] = None, **bounds_arguments):
    """One phase of the divide-and-conquer loop on one GPU, as far as this package goes: for every pile (a cells x genes
    UMI matrix) the k-NN graph (``compute_pile_knn_graph``'s steps, one pile after the other: each uses the whole GPU
    for ~10 ms) and then, for all piles side by side, the seeds and the first partition optimisation on the graphs as
    they lie on the device (``KnnGraph.seed_partition``: one SM, one host thread and stream per pile).  Returns a list of
    ``(community_of_nodes, score, seeds_count, k)``.  The graphs stay on the device in between (12 bytes per edge)."""
    bounds = metacell_bounds(**bounds_arguments)
    graphs = []
    try:
        for umis in piles:
            graphs.append(pile_knn_graph_on_device(umis, features=features, random_seed=random_seed))
        if not graphs:
            return []

        def candidates(entry):
            graph, cell_umis = entry
            wanted = seeds_count_for(float(sum(cell_umis)), graph.shape[0], bounds)
            communities, score, seeds = graph.seed_partition(cell_umis, bounds, wanted, random_seed=random_seed)
            return communities, score, seeds, graph.k

        return list(worker_pool(_side_by_side(workers)).map(candidates, graphs))
    finally:
        for graph, _ in graphs:
            graph.close()


def compute_direct_metacells(umis: sp.csr_matrix, *, features="select", random_seed: int = 123456,
                             must_complete_cover: bool = False, candidates_arguments: Optional[dict] = None,
                             deviants_arguments: Optional[dict] = None, dissolve_arguments: Optional[dict] = None,
                             noisy_genes_mask: Optional[np.ndarray] = None) -> dict:
    """One pile from its UMIs to its metacells, as ``compute_direct_metacells`` walks it (pipeline/direct.py:197-338):
    downsample -> feature genes -> log fractions -> similarity -> k-NN graph -> ``compute_candidate_metacells`` ->
    ``find_deviant_cells`` -> ``dissolve_metacells`` (the last two unless ``must_complete_cover``).  The pile is uploaded
    once; the graph, the candidates session and the dense matrices of the deviants never leave the device.  Returns
    ``metacell`` (int32 per cell, -1 = outlier), ``candidate``, ``deviant``, ``dissolved``, ``k`` and the graph's ``edges``."""
    from .candidates import compute_candidate_metacells
    from .deviants import find_deviant_cells
    from .dissolve import dissolve_metacells

    cell_umis = np.asarray(umis.sum(axis=1)).reshape(-1).astype("float32")
    samples = int(round(min(max(tl.DOWNSAMPLE_MIN_SAMPLES, np.quantile(cell_umis, tl.DOWNSAMPLE_MIN_CELL_QUANTILE)),
                            np.quantile(cell_umis, tl.DOWNSAMPLE_MAX_CELL_QUANTILE))))
    k = knn_k_of_pile(umis.shape[0], cell_umis)
    with Pile(umis, max_cell_total=float(cell_umis.max(initial=0))) as pile:
        _start_selection(pile, features)
        pile.downsample(samples, random_seed)
        graph = pile.knn_graph(_feature_genes_of(pile, features), k=k)
        try:
            candidate = compute_candidate_metacells(graph, cell_umis, random_seed=random_seed,
                                                    must_complete_cover=must_complete_cover, **(candidates_arguments or {}))
            result = {"candidate": candidate, "k": graph.k, "edges": graph.nnz}
        finally:
            graph.close()
        if must_complete_cover:  # direct.py:307-309
            assert candidate.min() == 0
            result.update(metacell=candidate, deviant=np.zeros(len(candidate), dtype="bool"),
                          dissolved=np.zeros(len(candidate), dtype="bool"))
            return result
        deviant = find_deviant_cells(pile, candidate, total_umis_per_cell=cell_umis, noisy_genes_mask=noisy_genes_mask,
                                     **(deviants_arguments or {}))
        metacell, dissolved = dissolve_metacells(pile, candidate, deviant, cell_umis, **(dissolve_arguments or {}))
    result.update(metacell=metacell, deviant=deviant, dissolved=dissolved)
    return result
