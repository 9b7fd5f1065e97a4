"""Host-side mirror of the hot-path functions of ``metacells.utilities.computation`` (same names,
argument meaning and results), on numpy / scipy.sparse data.  Every function is glue around
``metacells_b200.extensions`` exactly as the reference's functions are glue around
``metacells.extensions``; none of them computes on the CPU what the reference computes natively.
"""
from __future__ import annotations

from typing import Optional

import numpy as np
import scipy.sparse as sp

from . import extensions as xt

__all__ = [
    "downsample_matrix", "downsample_vector", "corrcoef", "cross_corrcoef_rows", "pairs_corrcoef_rows", "log_data",
    "rank_per", "rank_matrix_by_layout", "top_per", "prune_per", "relayout_compressed", "sort_compressed_indices",
]


def _as_csr_f32(matrix: sp.spmatrix) -> sp.csr_matrix:
    matrix = sp.csr_matrix(matrix)
    if matrix.data.dtype != np.float32:
        matrix = matrix.astype("float32")
    if matrix.indices.dtype != np.int32:
        matrix.indices = matrix.indices.astype("int32")
    if matrix.indptr.dtype != np.int32:
        matrix.indptr = matrix.indptr.astype("int32")
    return matrix


def downsample_matrix(matrix, *, per: str = "row", samples, eliminate_zeros: bool = True, random_seed: int = 0):
    """computation.py:746-901 for ``per="row"`` on CSR or row-major dense float32 data."""
    if per != "row":
        raise NotImplementedError("only per='row' is on the hot path (tools/downsample.py:99)")
    if sp.issparse(matrix):
        csr = _as_csr_f32(matrix)
        rows = csr.shape[0]
        samples_array = np.full(rows, samples, dtype="int32") if np.isscalar(samples) else np.asarray(samples, dtype="int32")
        output = np.empty_like(csr.data)
        xt.downsample_compressed_float32_t_int32_t_float32_t(csr.data, csr.indptr, output, samples_array, random_seed)
        result = sp.csr_matrix((output, csr.indices.copy(), csr.indptr.copy()), shape=csr.shape)
        if eliminate_zeros:  # computation.py:892-899
            result.eliminate_zeros()
            result.sort_indices()
        return result
    dense = np.ascontiguousarray(matrix, dtype="float32")
    samples_array = np.full(dense.shape[0], samples, dtype="int32") if np.isscalar(samples) else np.asarray(samples, dtype="int32")
    output = np.empty_like(dense)
    xt.downsample_dense_float32_t_float32_t(dense, output, samples_array, random_seed)
    return output


def downsample_vector(vector, samples: int, *, output: Optional[np.ndarray] = None, random_seed: int = 0):
    """computation.py:905-951."""
    source = np.ascontiguousarray(vector, dtype="float32")
    result = np.empty_like(source)
    xt.downsample_array_float32_t_float32_t(source, result, samples, random_seed)
    if output is not None:
        output[:] = result
        return output
    return result


def log_data(dense: np.ndarray, *, base: Optional[float] = 2, normalization: float = 0) -> np.ndarray:
    """computation.py:679-741 (dense float32, elementwise; numpy on the host as in the reference)."""
    result = np.array(dense, dtype="float32", copy=True)
    if normalization > 0:
        result += normalization
    if base == 2:
        np.log2(result, out=result)
    else:
        np.log(result, out=result)
        if base is not None:
            result /= np.log(base)
    return result


def corrcoef(matrix: np.ndarray, *, per: str = "row", reproducible: bool = True) -> np.ndarray:
    """computation.py:385-467 (the reproducible path: correlate_dense + NaN -> 0 + unit diagonal)."""
    del reproducible  # the GPU path is always reproducible
    dense = np.asarray(matrix)
    if per == "column":
        dense = dense.T
    if dense.dtype not in (np.float32, np.float64):
        dense = dense.astype("float32")
    if dense.shape[1] > 1 and dense.strides[1] != dense.itemsize:
        dense = np.ascontiguousarray(dense)
    result = np.empty((dense.shape[0], dense.shape[0]), dtype="float32")
    getattr(xt, f"correlate_dense_{dense.dtype}_t")(dense, result)
    result[np.isnan(result)] = 0.0  # computation.py:465 (already done on the device)
    np.fill_diagonal(result, 1.0)
    return result


def cross_corrcoef_rows(first: np.ndarray, second: np.ndarray, *, reproducible: bool = True) -> np.ndarray:
    """computation.py:471-530."""
    del reproducible
    dtype = "float64" if first.dtype == np.float64 and second.dtype == np.float64 else "float32"
    first = np.ascontiguousarray(first, dtype=dtype)
    second = np.ascontiguousarray(second, dtype=dtype)
    result = np.empty((first.shape[0], second.shape[0]), dtype="float32")
    getattr(xt, f"cross_correlate_dense_{dtype}_t")(first, second, result)
    return result


def pairs_corrcoef_rows(first: np.ndarray, second: np.ndarray, *, reproducible: bool = True) -> np.ndarray:
    """computation.py:534-575."""
    del reproducible
    dtype = "float64" if first.dtype == np.float64 and second.dtype == np.float64 else "float32"
    first = np.ascontiguousarray(first, dtype=dtype)
    second = np.ascontiguousarray(second, dtype=dtype)
    result = np.empty(first.shape[0], dtype="float32")
    getattr(xt, f"pairs_correlate_dense_{dtype}_t")(first, second, result)
    return result


def rank_per(matrix: np.ndarray, rank: int, *, per: str = "row") -> np.ndarray:
    """computation.py:1272-1294 (dense)."""
    dense = np.asarray(matrix, dtype="float32")
    if per == "column":
        dense = dense.T
    dense = np.ascontiguousarray(dense)
    output = np.empty(dense.shape[0], dtype="float32")
    xt.rank_rows_float32_t(dense, output, rank)
    return output


def rank_matrix_by_layout(matrix: np.ndarray, ascending: bool) -> np.ndarray:
    """computation.py:1691-1703: replaces the row-major matrix, in place, by its per-row ranks."""
    if matrix.dtype != np.float32:
        raise TypeError("rank_matrix_by_layout: float32 only on this path")
    xt.rank_matrix_float32_t(matrix, ascending)
    return matrix


def top_per(matrix: np.ndarray, top: int, *, per: str = "row", ranks: bool = False) -> sp.csr_matrix:
    """computation.py:1299-1352."""
    dense = np.asarray(matrix, dtype="float32")
    if per == "column":
        dense = dense.T
    dense = np.ascontiguousarray(dense)
    rows, columns = dense.shape
    if not 0 < top < columns:
        raise ValueError("top must be in 1 .. columns - 1")
    indptr = np.arange(rows + 1, dtype="int64") * top  # computation.py:1324
    indices = np.empty(rows * top, dtype="int32")
    data = np.empty(rows * top, dtype="float32")
    xt.collect_top_float32_t(top, dense, indices, data, ranks)
    result = sp.csr_matrix((data, indices, indptr), shape=dense.shape)
    result.has_sorted_indices = True
    return result.T.tocsc() if per == "column" else result


def prune_per(compressed, top: int):
    """computation.py:1356-1393: keep the ``top`` largest values of every band of a CSR / CSC matrix."""
    layout = compressed.getformat()
    if layout not in ("csr", "csc"):
        raise ValueError("prune_per needs a CSR or CSC matrix")
    data = np.ascontiguousarray(compressed.data, dtype="float32")
    indices = np.ascontiguousarray(compressed.indices, dtype="int32")
    indptr = np.ascontiguousarray(compressed.indptr, dtype="int32")
    bands = len(indptr) - 1
    capacity = max(min(bands * top, len(data)), 1)
    out_data = np.empty(capacity, dtype="float32")
    out_indices = np.empty(capacity, dtype="int32")
    out_indptr = np.empty(bands + 1, dtype="int32")
    xt.collect_pruned_float32_t_int32_t_int32_t(top, data, indices, indptr, out_data, out_indices, out_indptr)
    nnz = int(out_indptr[-1])
    constructor = sp.csr_matrix if layout == "csr" else sp.csc_matrix
    return constructor((out_data[:nnz], out_indices[:nnz], out_indptr), shape=compressed.shape)


def relayout_compressed(compressed):
    """computation.py:251-303 (_relayout_compressed): CSR -> CSC or CSC -> CSR."""
    layout = compressed.getformat()
    axis = ("csr", "csc").index(layout)
    data = np.ascontiguousarray(compressed.data, dtype="float32")
    indices = np.ascontiguousarray(compressed.indices, dtype="int32")
    indptr = np.ascontiguousarray(compressed.indptr, dtype="int32")
    out_bands = compressed.shape[1 - axis]
    counts = np.bincount(indices, minlength=out_bands)
    out_indptr = np.empty(out_bands + 1, dtype="int32")
    out_indptr[0:2] = 0
    np.cumsum(counts[:-1], out=out_indptr[2:])
    out_data = np.empty_like(data)
    out_indices = np.empty_like(indices)
    xt.collect_compressed_float32_t_int32_t_int32_t(data, indices, indptr, out_data, out_indices, out_indptr[1:])
    assert out_indptr[-1] == indptr[-1]
    constructor = (sp.csr_matrix, sp.csc_matrix)[1 - axis]
    result = constructor((out_data, out_indices, out_indptr), shape=compressed.shape)
    sort_compressed_indices(result, force=True)
    return result


def sort_compressed_indices(matrix, force: bool = False) -> None:
    """computation.py:306-330."""
    if matrix.has_sorted_indices and not force:
        return
    if matrix.data.dtype != np.float32 or matrix.indices.dtype != np.int32 or matrix.indptr.dtype != np.int32:
        raise TypeError("sort_compressed_indices: float32 / int32 / int32 only on this path")
    xt.sort_compressed_indices_float32_t_int32_t_int32_t(matrix.data, matrix.indices, matrix.indptr, matrix.shape[1])
    matrix.has_sorted_indices = True
