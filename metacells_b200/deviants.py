"""``find_deviant_cells`` with the (default) ``gaps`` policy (tools/deviants.py:31-137, :241-356) for a pile on the GPU.

The reference makes three dense cells x genes float32 matrices on the host (UMIs, fractions, log fractions: 1.7 GB each for
a 16 000 x 27 000 pile) and calls ``xt.compute_cell_gaps`` on them once per round.  Here the two matrices that call reads
are made on the device from the pile's sparse UMIs and stay there for all rounds (``mcb_pile_gap_inputs_create``); every
round sends the deviant / active flags and the per-gene thresholds and gets the per-cell gaps back
(``mcb_gap_inputs_cell_gaps``).  The values are the reference's bit for bit: the division and the addition are IEEE float32
operations, and numpy's float32 ``log2`` is taken by numpy itself, of one small table per cell (the counts 0 .. 255) plus
the few entries with larger counts.
"""
from __future__ import annotations

import ctypes
from typing import Optional

import numpy as np

from ._lib import check, lib

__all__ = ["GapInputs", "find_deviant_cells", "regularization_of_cells"]

# metacells/parameters.py:34, :43, :677-758
DEVIANTS_MIN_GENE_FOLD_FACTOR = 3.0
DEVIANTS_MIN_NOISY_GENE_FOLD_FACTOR = 2.0
DEVIANTS_MIN_COMPARE_UMIS = 8
DEVIANTS_GAP_SKIP_CELLS = 1
DEVIANTS_MAX_CELL_FRACTION = 0.25
DEVIANTS_MAX_GAP_CELLS_COUNT = 3
DEVIANTS_MAX_GAP_CELLS_FRACTION = 0.1
DEVIANT_CELLS_REGULARIZATION_QUANTILE = 0.25
LOG_TABLE_COUNTS = 256


def _pointer(array: np.ndarray) -> ctypes.c_void_p:
    return ctypes.c_void_p(array.ctypes.data)


def regularization_of_cells(total_umis_per_cell: np.ndarray, candidate_per_cell: np.ndarray, candidates_count: int,
                            cells_regularization_quantile: float) -> np.ndarray:
    """tools/deviants.py:283-291: one over the larger of the cell's total UMIs and the candidate's quantile of them."""
    regularization_per_cell = np.empty(len(total_umis_per_cell), dtype="float32")
    for candidate in range(candidates_count):
        mask = candidate_per_cell == candidate
        totals = total_umis_per_cell[mask]
        floor = max(1.0, np.quantile(totals, cells_regularization_quantile))
        regularization_per_cell[mask] = 1.0 / np.maximum(totals, floor)
    return regularization_per_cell


class GapInputs:
    """The dense UMIs and log fractions of a pile's cells on the GPU (``mcb_gap_inputs``), as ``compute_cell_gaps`` reads
    them (tools/deviants.py:270-292)."""

    def __init__(self, pile, total_umis_per_cell: np.ndarray, regularization_per_cell: np.ndarray, *,
                 table_counts: int = LOG_TABLE_COUNTS):
        total_umis_per_cell = np.ascontiguousarray(total_umis_per_cell, dtype="float32")
        regularization_per_cell = np.ascontiguousarray(regularization_per_cell, dtype="float32")
        self.shape = pile.shape
        assert len(total_umis_per_cell) == len(regularization_per_cell) == self.shape[0]
        counts = np.arange(table_counts, dtype="float32")
        with np.errstate(divide="ignore", invalid="ignore"):  # a cell without UMIs: the reference takes log2(nan) as well
            table = np.log2(counts[np.newaxis, :] / total_umis_per_cell[:, np.newaxis] + regularization_per_cell[:, np.newaxis])
        table = np.ascontiguousarray(table, dtype="float32")
        self._handle = ctypes.c_void_p()
        function = lib().mcb_pile_gap_inputs_create
        function.restype = ctypes.c_int
        check(function(pile._handle, _pointer(total_umis_per_cell), _pointer(regularization_per_cell), _pointer(table),
                       ctypes.c_int64(table_counts), ctypes.byref(self._handle)))
        count_of = lib().mcb_gap_inputs_exceptions
        count_of.restype = ctypes.c_int64
        self.exceptions = int(count_of(self._handle))
        if self.exceptions:
            values = np.empty(self.exceptions, dtype="float32")
            function = lib().mcb_gap_inputs_read_exceptions
            function.restype = ctypes.c_int
            check(function(self._handle, _pointer(values)))
            logs = np.log2(values)
            function = lib().mcb_gap_inputs_patch
            function.restype = ctypes.c_int
            check(function(self._handle, _pointer(logs)))

    def close(self) -> None:
        if getattr(self, "_handle", None):
            function = lib().mcb_gap_inputs_free
            function.restype = None
            function(self._handle)
            self._handle = None

    def __del__(self):
        self.close()

    def __enter__(self):
        return self

    def __exit__(self, *_):
        self.close()

    def read(self):
        """``(umis_per_gene_per_cell, log_fraction_per_gene_per_cell)`` as host arrays (tests, small piles)."""
        umis, logs = np.empty(self.shape, dtype="float32"), np.empty(self.shape, dtype="float32")
        function = lib().mcb_gap_inputs_read
        function.restype = ctypes.c_int
        check(function(self._handle, _pointer(umis), _pointer(logs)))
        return umis, logs

    def device_arrays(self):
        """``(umis, log_fraction)``: device addresses of the two float32 [cells, genes] matrices."""
        umis, logs = ctypes.c_void_p(), ctypes.c_void_p()
        function = lib().mcb_gap_inputs_device_arrays
        function.restype = ctypes.c_int
        check(function(self._handle, ctypes.byref(umis), ctypes.byref(logs)))
        return umis.value, logs.value

    def compute_cell_gaps(self, candidate_per_cell: np.ndarray, deviant_per_cell: np.ndarray, active_per_cell: np.ndarray,
                          min_gap_per_gene: np.ndarray, candidates_count: int, gap_skip_cells: int, max_gap_cells_count: int,
                          max_gap_cells_fraction: float, min_compare_umis: float, max_gap_per_cell: np.ndarray) -> None:
        """``xt.compute_cell_gaps`` (gaps.cpp:6-186) on the resident matrices; ``max_gap_per_cell`` is updated in place."""
        assert candidate_per_cell.dtype == np.int32 and max_gap_per_cell.dtype == np.float32 and min_gap_per_gene.dtype == np.float32
        deviant = np.ascontiguousarray(deviant_per_cell, dtype="bool")
        active = np.ascontiguousarray(active_per_cell, dtype="bool")
        function = lib().mcb_gap_inputs_cell_gaps
        function.restype = ctypes.c_int
        check(function(self._handle, _pointer(candidate_per_cell), _pointer(deviant), _pointer(active), _pointer(min_gap_per_gene),
                       ctypes.c_int64(int(candidates_count)), ctypes.c_int64(int(gap_skip_cells)),
                       ctypes.c_int64(int(max_gap_cells_count)), ctypes.c_float(max_gap_cells_fraction),
                       ctypes.c_float(min_compare_umis), _pointer(max_gap_per_cell)))


def find_deviant_cells(pile, candidates: np.ndarray, *, total_umis_per_cell: Optional[np.ndarray] = None,
                       noisy_genes_mask: Optional[np.ndarray] = None, min_gene_fold_factor: float = DEVIANTS_MIN_GENE_FOLD_FACTOR,
                       min_compare_umis: int = DEVIANTS_MIN_COMPARE_UMIS, gap_skip_cells: int = DEVIANTS_GAP_SKIP_CELLS,
                       min_noisy_gene_fold_factor: float = DEVIANTS_MIN_NOISY_GENE_FOLD_FACTOR,
                       max_cell_fraction: Optional[float] = DEVIANTS_MAX_CELL_FRACTION,
                       max_gap_cells_count: int = DEVIANTS_MAX_GAP_CELLS_COUNT,
                       max_gap_cells_fraction: float = DEVIANTS_MAX_GAP_CELLS_FRACTION,
                       cells_regularization_quantile: float = DEVIANT_CELLS_REGULARIZATION_QUANTILE, policy: str = "gaps",
                       trace: Optional[list] = None) -> np.ndarray:
    """``find_deviant_cells`` (tools/deviants.py:31-137 -> ``_gaps_deviant_cells``, :241-356) for a ``Pile`` whose UMIs are
    on the GPU: the boolean mask of the cells that stand apart from their candidate metacell by a gap in some gene.
    ``candidates``: the candidate of every cell; ``total_umis_per_cell``: the row sums of the pile (computed if absent);
    ``noisy_genes_mask``: the ``noisy_gene`` annotation.  Only the reference's default policy is provided."""
    if policy != "gaps":
        raise NotImplementedError("only the default 'gaps' policy of find_deviant_cells is provided")
    assert 0 <= gap_skip_cells <= 2
    gap_skip_cells += 1  # :100
    max_cell_fraction = max_cell_fraction or 1.0
    assert min_gene_fold_factor > 0 and 0 < max_cell_fraction < 1
    cells_count, genes_count = pile.shape
    if cells_count <= max_gap_cells_count:  # :260-261
        return np.zeros(cells_count, dtype="bool")

    min_gap_per_gene = np.full(genes_count, min_gene_fold_factor, dtype="float32")
    if noisy_genes_mask is not None:
        min_gap_per_gene[np.asarray(noisy_genes_mask, dtype=bool)] += min_noisy_gene_fold_factor
    candidate_per_cell = np.ascontiguousarray(candidates).astype("int32")
    candidates_count = int(np.max(candidate_per_cell)) + 1
    if total_umis_per_cell is None:
        total_umis_per_cell = pile.cell_totals()
    total_umis_per_cell = np.ascontiguousarray(total_umis_per_cell, dtype="float32")
    regularization_per_cell = regularization_of_cells(total_umis_per_cell, candidate_per_cell, candidates_count,
                                                      cells_regularization_quantile)
    max_gap_per_cell = np.empty(cells_count, dtype="float32")

    with GapInputs(pile, total_umis_per_cell, regularization_per_cell) as inputs:
        while True:  # :294-356
            new_deviant_per_cell = np.ones(cells_count, dtype="bool")
            deviant_per_cell = np.zeros(cells_count, dtype="bool")
            deviants_fraction = 0.0
            while deviants_fraction <= max_cell_fraction:
                max_gap_per_cell[:] = 0.0
                inputs.compute_cell_gaps(candidate_per_cell, deviant_per_cell, new_deviant_per_cell, min_gap_per_gene,
                                         candidates_count, gap_skip_cells, max_gap_cells_count, max_gap_cells_fraction,
                                         float(min_compare_umis), max_gap_per_cell)
                if trace is not None:
                    trace.append(max_gap_per_cell.copy())
                large_gap_count = np.sum(max_gap_per_cell > 0.0)
                remaining_cell_fraction = 1 - max(max_cell_fraction - deviants_fraction, 0.0)
                if large_gap_count / cells_count <= remaining_cell_fraction:
                    gap_threshold = 0.0
                else:
                    gap_threshold = np.quantile(max_gap_per_cell, remaining_cell_fraction)
                new_deviant_per_cell = (max_gap_per_cell > gap_threshold) & ~deviant_per_cell
                new_deviants_count = np.sum(new_deviant_per_cell)
                if new_deviants_count == 0:
                    break
                deviants_fraction += new_deviants_count / cells_count
                deviant_per_cell = deviant_per_cell | new_deviant_per_cell
            if deviants_fraction <= max_cell_fraction:
                return deviant_per_cell
            min_gap_per_gene += 1 / 8
            min_gene_fold_factor += 1 / 8
