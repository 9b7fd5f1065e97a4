"""Feature-gene selection of a pile (``extract_selected_data``, pipeline/select.py:30-220, over tools/high.py:29-135 and
:246-303), with the O(non-zeros) part on the GPU.

The reference densifies the downsampled cells x genes matrix column-major (1.7 GB for a 16 000 x 27 000 pile) to rank every
gene's column (tools/high.py:121-123 -> computation.py:1285-1293) and sums the CSC matrix twice.  Here the downsample kernel
adds up, per gene, the total, the exact sum of squares and the number of cells at or above ``min_gene_top3`` where it writes
every count (``mcb_pile_want_gene_statistics``); what is left is arithmetic on vectors of one entry per gene, done in numpy
with the reference's own operations and dtypes (float32 sums, float64 variance, float32 log2), plus the 100-gene sliding
median on the GPU (the median of an odd window is one of its elements: ``mcb_host_sliding_median_f32`` against 15 ms of
``pandas`` rolling median), so the masks are the reference's bit for bit:

* a float32 sum of integers is exact, in any order, while it stays below 2^24; the few genes whose sum of squares (or total)
  passes 2^24 are the only ones where the order of numpy's pairwise summation shows, and their columns are fetched
  (``mcb_pile_gene_columns``) and re-summed by ``np.add.reduceat`` itself, as ``scipy.sparse``'s ``sum(axis=0)`` of a CSC
  matrix does;
* ``rank_per(column, rank=max(n_obs - topN - 1, 1)) >= min_gene_topN`` holds exactly when at least ``n_obs - rank`` cells hold
  ``>= min_gene_topN`` UMIs of the gene: a counter per gene instead of a selection per column.
"""
from __future__ import annotations

from typing import Optional, Tuple

import numpy as np

__all__ = ["gene_sums", "relative_variance_of_genes", "select_feature_genes", "SELECT_MIN_GENE_RELATIVE_VARIANCE",
           "SELECT_MIN_GENE_TOTAL", "SELECT_MIN_GENE_TOP3", "SELECT_MIN_GENES", "RELATIVE_VARIANCE_WINDOW_SIZE"]

# metacells/parameters.py:456-482, :79
SELECT_MIN_GENE_RELATIVE_VARIANCE = 0.1
SELECT_MIN_GENE_TOTAL = 50
SELECT_MIN_GENE_TOP3 = 4
SELECT_MIN_GENES = 100
RELATIVE_VARIANCE_WINDOW_SIZE = 100

_EXACT_BELOW = float(1 << 24)


def gene_sums(pile) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
    """``(sum_per, sum_squared_per, top_counts)`` of the downsampled pile's genes: the two float32 vectors
    ``scipy.sparse``'s ``sum(axis=0)`` gives for the column-major matrix and its square (computation.py:1224-1268), and the
    number of cells holding at least the threshold given to ``want_gene_statistics``."""
    totals, squares, top_counts = pile.gene_statistics()
    sums = totals.copy()
    sums_squared = squares.astype("float32")  # exact wherever the integer is below 2^24
    inexact = np.flatnonzero((squares >= (1 << 24)) | (totals >= _EXACT_BELOW)).astype("int32")
    if len(inexact):
        values, sizes = pile.gene_columns(inexact)
        for slot, gene in enumerate(inexact):
            column = values[slot, :sizes[slot]]
            # scipy's _minor_reduce: ufunc.reduceat over the column's stored values (numpy's pairwise float32 sum)
            sums[gene] = np.add.reduceat(column, [0])[0]
            sums_squared[gene] = np.add.reduceat(column * column, [0])[0]
    return sums, sums_squared, top_counts


def _device_window_medians(values: np.ndarray, window_size: int) -> np.ndarray:
    """``pd.Series(values).rolling(window_size).median()`` without its leading NaNs, on the GPU
    (``mcb_host_sliding_median_f32``): the median of an odd window is one of its elements."""
    import ctypes

    from ._lib import check, lib

    values = np.ascontiguousarray(values, dtype="float32")
    medians = np.empty(len(values) - window_size + 1, dtype="float32")
    function = lib().mcb_host_sliding_median_f32
    function.restype = ctypes.c_int
    check(function(ctypes.c_void_p(values.ctypes.data), ctypes.c_int64(len(values)), ctypes.c_int64(window_size),
                   ctypes.c_void_p(medians.ctypes.data)))
    return medians.astype("float64")


def _device_ordered_medians(values: np.ndarray, order: np.ndarray, window_size: int) -> np.ndarray:
    """The medians of ``_sliding_median`` for a given ``order = argsort(order_by)``, in entry order (float64): the gathers,
    the padding of the end windows and the way back all on the GPU (``mcb_host_sliding_median_ordered_f32``)."""
    import ctypes

    from ._lib import check, lib

    values = np.ascontiguousarray(values, dtype="float32")
    order = np.ascontiguousarray(order, dtype="int64")
    medians = np.empty(len(values), dtype="float64")
    function = lib().mcb_host_sliding_median_ordered_f32
    function.restype = ctypes.c_int
    check(function(ctypes.c_void_p(values.ctypes.data), ctypes.c_void_p(order.ctypes.data), ctypes.c_int64(len(values)),
                   ctypes.c_int64(window_size), ctypes.c_void_p(medians.ctypes.data)))
    return medians


_sorter = None


def _argsort_meanwhile(order_by: np.ndarray):
    """``np.argsort(order_by)`` on a helper thread (numpy sorts without the interpreter lock): the variance arithmetic of
    the caller runs beside it.  The order of equal keys is numpy's own and part of the result, so numpy does the sort."""
    global _sorter
    if _sorter is None:
        from concurrent.futures import ThreadPoolExecutor

        _sorter = ThreadPoolExecutor(max_workers=4, thread_name_prefix="mcb-argsort")
    return _sorter.submit(np.argsort, order_by)


def _sliding_median(values: np.ndarray, order_by: np.ndarray, window_size: int, window_medians=None, order=None) -> np.ndarray:
    """``sliding_window_function(values, function="median", window_size=..., order_by=...)`` (computation.py:1836-1897):
    the median of the ``window_size`` entries nearest in ``order_by``, the windows at both ends padded with the entries
    the reference pads them with.  ``window_medians(values, window_size)``: the medians of all whole windows (default: on
    the GPU)."""
    if window_size % 2 == 0:
        window_size += 1
    half = (window_size - 1) // 2
    count = len(values)
    if window_size >= count:
        return np.full(count, np.median(values))
    if order is None:
        order = np.argsort(order_by)
    elif hasattr(order, "result"):
        order = order.result()
    if window_medians is None and window_size <= 127:
        return _device_ordered_medians(values, order, window_size)
    back = np.empty(count, dtype=order.dtype)  # == np.argsort(order): the inverse of a permutation, by scatter
    back[order] = np.arange(count, dtype=order.dtype)
    padded = np.concatenate([order[window_size - half:window_size], order, order[count - window_size:count - window_size + half]])
    medians = (window_medians or _device_window_medians)(values[padded], window_size)
    assert medians.size == count
    return medians[back]


def relative_variance_of_genes(sums: np.ndarray, sums_squared: np.ndarray, cells: int,
                               window_size: int = RELATIVE_VARIANCE_WINDOW_SIZE, window_medians=None) -> np.ndarray:
    """``relative_variance_per(column_major, per="column")`` (computation.py:1571-1595) from the per-gene sums: variance
    (:1526-1546, float64 from the squared float32 sum), normalized variance (:1549-1568, float32), its ``log2`` minus the
    median over the 100 genes of the nearest mean expression."""
    assert sums.dtype == np.float32 and sums_squared.dtype == np.float32
    variance = np.square(sums).astype(float)
    variance /= -cells
    variance += sums_squared
    variance /= cells
    variance[variance < 0] = 0
    mean = sums / cells
    order = _argsort_meanwhile(mean) if len(mean) > max(window_size, 4096) else None
    silent = mean == 0
    normalized = np.reciprocal(mean, where=~silent)
    normalized[silent] = 0
    normalized *= variance
    normalized[silent] = 1.0
    with np.errstate(divide="ignore"):
        np.log2(normalized, out=normalized)
    return normalized - _sliding_median(normalized, mean, window_size, window_medians, order)


def select_feature_genes(pile, *, min_gene_relative_variance: Optional[float] = SELECT_MIN_GENE_RELATIVE_VARIANCE,
                         min_genes: int = SELECT_MIN_GENES, min_gene_total: Optional[int] = SELECT_MIN_GENE_TOTAL,
                         min_gene_top3: Optional[int] = SELECT_MIN_GENE_TOP3,
                         lateral_genes_mask: Optional[np.ndarray] = None, window_medians=None) -> np.ndarray:
    """The ``selected_gene`` mask of ``extract_selected_data`` (pipeline/select.py:145-220) for a downsampled ``Pile`` whose
    statistics were asked for (``pile.want_gene_statistics(min_gene_top3)`` before ``pile.downsample``): genes with a high
    top-3 value, a high total and a high relative variance that are not lateral; when fewer than ``min_genes`` pass, the
    relative-variance bar is lowered in steps of 1/8 and bisected four times (:176-209), or dropped when even the other
    masks leave too few."""
    assert min_genes > 0
    cells, genes = pile.shape
    sums, sums_squared, top_counts = gene_sums(pile)
    masks = []
    if min_gene_top3 is not None:
        assert pile.gene_top_threshold == float(min_gene_top3), "the pile counted cells against another top-3 threshold"
        rank = max(cells - 3 - 1, 1)  # tools/high.py:122
        masks.append(top_counts >= cells - rank)
    if min_gene_total is not None:
        masks.append(sums >= min_gene_total)  # tools/high.py:73-74
    relative = None
    if min_gene_relative_variance is not None:
        relative = relative_variance_of_genes(sums, sums_squared, cells, window_medians=window_medians)
        masks.append(relative >= min_gene_relative_variance)
    fixed = [] if lateral_genes_mask is None else [~np.asarray(lateral_genes_mask, dtype=bool)]

    def passing(some) -> np.ndarray:
        return np.logical_and.reduce(some + fixed) if some or fixed else np.ones(genes, dtype=bool)

    selected = passing(masks)
    if selected.sum() >= min_genes:
        return selected
    if min_gene_relative_variance is None:
        raise ValueError(f"Failed to select {min_genes} genes")
    others = passing(masks[:-1])
    if others.sum() <= min_genes:  # :180-181: the relative variance is given up
        return others
    bar = min_gene_relative_variance
    while True:  # :184-195
        high, bar = bar, bar - 1 / 8
        kept = passing(masks[:-1] + [relative >= bar])
        if kept.sum() >= min_genes:
            break
    last_mask, last_valid = kept, True
    for _ in range(4):  # :197-209
        middle = (high + bar) / 2.0
        last_mask = passing(masks[:-1] + [relative >= middle])
        last_valid = bool(last_mask.sum() >= min_genes)
        if last_valid:
            bar, kept = middle, last_mask
        else:
            high = middle
    # the reference filters by the mask of its last find_high_relative_variance_genes call, unless that was too strict
    return last_mask if last_valid else kept
