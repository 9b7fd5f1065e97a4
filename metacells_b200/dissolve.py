"""``dissolve_metacells`` (tools/dissolve.py:24-223) for a pile on the GPU: the last step of a pile
(pipeline/direct.py:324-338).  Deviant cells become outliers; a candidate under the minimal size, or neither robust in
size or UMIs nor holding a gene whose fold against the whole pile is convincing, is dissolved into outliers.

The decisions are the reference's host arithmetic on one value per candidate or per gene; what touches the UMIs -- the
per-gene totals of the pile and of the few candidates that reach the fold test -- is summed on the device
(``mcb_pile_sum_groups``, exact integers) instead of from a column-major copy of the pile.
"""
from __future__ import annotations

import ctypes
from math import floor
from typing import Optional, Tuple

import numpy as np

from ._lib import check, lib

__all__ = ["dissolve_metacells", "compress_indices"]

# metacells/parameters.py:180, :767, :797
MIN_METACELL_SIZE = 12
TARGET_METACELL_SIZE = 48
TARGET_METACELL_UMIS = 160000
DISSOLVE_MIN_ROBUST_SIZE_FACTOR = 0.5
DISSOLVE_MIN_CONVINCING_GENE_FOLD_FACTOR = 3.0

_EXACT_BELOW = float(1 << 24)


def compress_indices(indices: np.ndarray) -> np.ndarray:
    """``ut.compress_indices`` (utilities/computation.py:1936-1946): consecutive group indices, -1 kept."""
    unique, consecutive = np.unique(indices, return_inverse=True)
    consecutive += min(unique[0], 0)
    return consecutive.astype(indices.dtype)


def _sum_groups(pile, group_of_cells: np.ndarray, groups: int) -> np.ndarray:
    """float64 [groups, genes]: the exact per-gene UMIs of the cells of every group."""
    group_of_cells = np.ascontiguousarray(group_of_cells, dtype="int32")
    sums = np.empty((groups, pile.shape[1]), dtype="float64")
    function = lib().mcb_pile_sum_groups
    function.restype = ctypes.c_int
    check(function(pile._handle, ctypes.c_void_p(group_of_cells.ctypes.data), ctypes.c_int64(groups),
                   ctypes.c_void_p(sums.ctypes.data)))
    return sums


def _float32_column_sums(pile, exact_totals: np.ndarray) -> np.ndarray:
    """``sum_per(data, per="column")`` of the pile (float32): an integer below 2^24 is what numpy's float32 sum gives in
    any order; the genes beyond get their column (stored values, in cell order) summed by numpy itself, as scipy's
    ``sum(axis=0)`` of a column-major matrix does."""
    totals = exact_totals.astype("float32")
    beyond = np.flatnonzero(exact_totals >= _EXACT_BELOW).astype("int32")
    if len(beyond):
        values, sizes = pile.gene_columns(beyond)
        for slot, gene in enumerate(beyond):
            totals[gene] = np.add.reduceat(values[slot, :sizes[slot]], [0])[0]
    return totals


def dissolve_metacells(pile, candidates: np.ndarray, deviants: Optional[np.ndarray], cell_umis: np.ndarray, *,
                       target_metacell_size: int = TARGET_METACELL_SIZE, min_metacell_size: int = MIN_METACELL_SIZE,
                       target_metacell_umis: int = TARGET_METACELL_UMIS,
                       min_robust_size_factor: float = DISSOLVE_MIN_ROBUST_SIZE_FACTOR,
                       min_convincing_gene_fold_factor: Optional[float] = DISSOLVE_MIN_CONVINCING_GENE_FOLD_FACTOR
                       ) -> Tuple[np.ndarray, np.ndarray]:
    """``(metacell_of_cells, dissolved_of_cells)`` as the reference sets them (tools/dissolve.py:81-132)."""
    assert cell_umis.dtype == np.float32
    cells_count = pile.shape[0]
    dissolved_of_cells = np.zeros(cells_count, dtype="bool")
    candidate_of_cells = np.copy(candidates)
    if deviants is not None:
        candidate_of_cells[np.asarray(deviants) > 0] = -1
    candidate_of_cells = compress_indices(candidate_of_cells)
    candidates_count = int(np.max(candidate_of_cells)) + 1
    min_robust_size = int(floor(target_metacell_size * min_robust_size_factor))
    min_robust_umis = int(floor(target_metacell_umis * min_robust_size_factor))

    # :147-184: what size and UMIs decide; the candidates left over go through the fold test
    members = [np.where(candidate_of_cells == candidate)[0] for candidate in range(candidates_count)]
    keep = np.ones(candidates_count, dtype="bool")
    undecided = []
    for candidate, cells in enumerate(members):
        if cells.size < min_metacell_size:
            keep[candidate] = False
        elif cells.size >= min_robust_size or np.sum(cell_umis[cells]) >= min_robust_umis:
            pass
        elif min_convincing_gene_fold_factor is not None:
            undecided.append(candidate)
    if undecided:  # :186-199
        group_of_cells = np.full(cells_count, len(undecided), dtype="int32")  # the last group: every other cell
        for group, candidate in enumerate(undecided):
            group_of_cells[members[candidate]] = group
        sums = _sum_groups(pile, group_of_cells, len(undecided) + 1)
        total_of_genes = _float32_column_sums(pile, sums.sum(axis=0))
        fraction_of_genes = total_of_genes / total_of_genes.sum()  # :101-102 (ut.fraction_per)
        for group, candidate in enumerate(undecided):
            data_of_genes = sums[group].astype("float32")  # far below 2^24: a candidate under min_robust_umis
            expected_of_genes = fraction_of_genes * np.sum(data_of_genes)
            expected_of_genes += 1
            data_of_genes += 1
            data_of_genes /= expected_of_genes
            np.log2(data_of_genes, out=data_of_genes)
            keep[candidate] = bool(np.any(np.abs(data_of_genes) >= min_convincing_gene_fold_factor))
    for candidate in np.flatnonzero(~keep):
        dissolved_of_cells[members[candidate]] = True
        candidate_of_cells[members[candidate]] = -1
    metacell_of_cells = compress_indices(candidate_of_cells) if not keep.all() else candidate_of_cells
    return metacell_of_cells, dissolved_of_cells
