"""Drop-in for the hot-path symbols of ``metacells.extensions`` (the reference's pybind11 module,
metacells/extensions.cpp:64-83).  Same function names (``<op>_<dtype>_t...``), same positional
arguments, same in-place output convention: the caller allocates every output with numpy and the
function fills it.  Each function forwards to the ``mcb_host_*`` entry point of libmcb200.so
(include/mcb200.h), which copies the buffers to the GPU, runs the sm_100a kernels and copies the
results back.  There is no CPU implementation here.

Differences from the reference, all supersets: argument errors raise ``TypeError`` /
``ValueError`` / ``McbError`` instead of printing and calling ``std::exit(1)``
(extensions.h:59-72), and a dtype mismatch is an error rather than a silent temporary copy
(pybind11 forcecast).  Only the instantiations the path uses exist.
"""
from __future__ import annotations

import numpy as np

from ._lib import McbError, call, call_f64, host_ptr, row_stride  # noqa: F401

__all__ = ["McbError", "set_threads_count"]


def set_threads_count(count: int) -> None:
    """extensions.cpp:13-18.  The GPU path has no host thread pool; accepted and ignored."""
    if int(count) < 0:
        raise ValueError("threads count must not be negative")


# ---- downsample.cpp ---------------------------------------------------------------------------


def downsample_compressed_float32_t_int32_t_float32_t(input_data, input_indptr, output, samples, random_seed):
    """downsample.cpp:171-192."""
    bands = len(input_indptr) - 1
    if len(samples) != bands:
        raise ValueError("samples must have one entry per band")
    if len(output) != len(input_data) or (bands >= 0 and len(input_data) != int(input_indptr[-1])):
        raise ValueError("data / output / indptr sizes disagree")
    call("mcb_host_downsample_compressed_f32_i32_f32", host_ptr(input_data, "float32", "input_data"),
         host_ptr(input_indptr, "int32", "input_indptr"), host_ptr(output, "float32", "output", True),
         host_ptr(samples, "int32", "samples"), ("i64", bands), ("u64", int(random_seed)))


def downsample_dense_float32_t_float32_t(input_matrix, output, samples, random_seed):
    """downsample.cpp:131-151."""
    if input_matrix.shape != output.shape or len(samples) != input_matrix.shape[0]:
        raise ValueError("input / output / samples shapes disagree")
    if not (input_matrix.flags.c_contiguous and output.flags.c_contiguous):
        raise ValueError("downsample_dense needs C-contiguous matrices")
    call("mcb_host_downsample_dense_f32_f32", host_ptr(input_matrix, "float32", "input"),
         host_ptr(output, "float32", "output", True), host_ptr(samples, "int32", "samples"),
         ("i64", input_matrix.shape[0]), ("i64", input_matrix.shape[1]), ("u64", int(random_seed)))


def downsample_array_float32_t_float32_t(input_array, output, samples, random_seed):
    """downsample.cpp:116-128."""
    if len(input_array) != len(output):
        raise ValueError("input / output sizes disagree")
    call("mcb_host_downsample_array_f32_f32", host_ptr(input_array, "float32", "input"),
         host_ptr(output, "float32", "output", True), ("i64", len(input_array)), ("i32", int(samples)),
         ("u64", int(random_seed)))


# ---- correlate.cpp ----------------------------------------------------------------------------


def _correlate_dense(input_matrix, output, suffix, dtype):
    rows, columns = input_matrix.shape
    if output.shape != (rows, rows) or not output.flags.c_contiguous:
        raise ValueError("output must be a C-contiguous [rows, rows] matrix")
    call(f"mcb_host_correlate_dense_{suffix}", host_ptr(input_matrix, dtype, "input"), ("i64", rows), ("i64", columns),
         ("i64", row_stride(input_matrix)), host_ptr(output, "float32", "output", True))


def correlate_dense_float32_t(input_matrix, output):
    """correlate.cpp:350-431."""
    _correlate_dense(input_matrix, output, "f32", "float32")


def correlate_dense_float64_t(input_matrix, output):
    """correlate.cpp:350-431 (float64 instantiation; output stays float32)."""
    _correlate_dense(input_matrix, output, "f64", "float64")


def _cross_correlate_dense(first, second, output, suffix, dtype):
    if first.shape[1] != second.shape[1]:
        raise ValueError("first and second must have the same number of columns")
    if output.shape != (first.shape[0], second.shape[0]) or not output.flags.c_contiguous:
        raise ValueError("output must be a C-contiguous [first rows, second rows] matrix")
    call(f"mcb_host_cross_correlate_dense_{suffix}", host_ptr(first, dtype, "first"), ("i64", first.shape[0]),
         ("i64", row_stride(first)), host_ptr(second, dtype, "second"), ("i64", second.shape[0]),
         ("i64", row_stride(second)), ("i64", first.shape[1]), host_ptr(output, "float32", "output", True))


def cross_correlate_dense_float32_t(first, second, output):
    """correlate.cpp:433-494."""
    _cross_correlate_dense(first, second, output, "f32", "float32")


def cross_correlate_dense_float64_t(first, second, output):
    """correlate.cpp:433-494."""
    _cross_correlate_dense(first, second, output, "f64", "float64")


def _pairs_correlate_dense(first, second, output, suffix, dtype):
    if first.shape != second.shape or len(output) != first.shape[0]:
        raise ValueError("first / second / output shapes disagree")
    call(f"mcb_host_pairs_correlate_dense_{suffix}", host_ptr(first, dtype, "first"), ("i64", row_stride(first)),
         host_ptr(second, dtype, "second"), ("i64", row_stride(second)), ("i64", first.shape[0]),
         ("i64", first.shape[1]), host_ptr(output, "float32", "output", True))


def pairs_correlate_dense_float32_t(first, second, output):
    """correlate.cpp:496-527."""
    _pairs_correlate_dense(first, second, output, "f32", "float32")


def pairs_correlate_dense_float64_t(first, second, output):
    """correlate.cpp:496-527."""
    _pairs_correlate_dense(first, second, output, "f64", "float64")


# ---- rank.cpp / top_per.cpp -------------------------------------------------------------------


def rank_matrix_float32_t(matrix, ascending):
    """rank.cpp:79-87: in place."""
    call("mcb_host_rank_matrix_f32", host_ptr(matrix, "float32", "matrix", True), ("i64", matrix.shape[0]),
         ("i64", matrix.shape[1]), ("i64", row_stride(matrix)), ("int", int(bool(ascending))))


def rank_rows_float32_t(input_matrix, output, rank):
    """rank.cpp:24-36."""
    if len(output) != input_matrix.shape[0]:
        raise ValueError("output must have one entry per row")
    call("mcb_host_rank_rows_f32", host_ptr(input_matrix, "float32", "input"), ("i64", input_matrix.shape[0]),
         ("i64", input_matrix.shape[1]), ("i64", row_stride(input_matrix)), host_ptr(output, "float32", "output", True),
         ("i64", int(rank)))


def collect_top_float32_t(degree, similarity, output_indices, output_data, ranks):
    """top_per.cpp:67-92."""
    rows = similarity.shape[0]
    if len(output_indices) != rows * degree or len(output_data) != rows * degree:
        raise ValueError("outputs must have rows * degree entries")
    call("mcb_host_collect_top_f32", ("i64", int(degree)), host_ptr(similarity, "float32", "similarity"), ("i64", rows),
         ("i64", similarity.shape[1]), ("i64", row_stride(similarity)), host_ptr(output_indices, "int32", "output_indices", True),
         host_ptr(output_data, "float32", "output_data", True), ("int", int(bool(ranks))))


# ---- prune_per.cpp / relayout.cpp -------------------------------------------------------------


def collect_pruned_float32_t_int32_t_int32_t(pruned_degree, input_data, input_indices, input_indptr, output_data,
                                             output_indices, output_indptr):
    """prune_per.cpp:56-105."""
    bands = len(input_indptr) - 1
    if len(output_indptr) != bands + 1:
        raise ValueError("output indptr must have bands + 1 entries")
    if len(input_data) != len(input_indices) or len(input_data) != int(input_indptr[-1]):
        raise ValueError("data / indices / indptr sizes disagree")
    needed = int(np.minimum(np.diff(input_indptr), pruned_degree).sum())
    if len(output_data) < needed or len(output_indices) < needed:
        raise ValueError("output arrays are too small")
    call("mcb_host_collect_pruned_f32_i32_i32", ("i64", int(pruned_degree)), host_ptr(input_data, "float32", "input_data"),
         host_ptr(input_indices, "int32", "input_indices"), host_ptr(input_indptr, "int32", "input_indptr"), ("i64", bands),
         host_ptr(output_data, "float32", "output_data", True), host_ptr(output_indices, "int32", "output_indices", True),
         host_ptr(output_indptr, "int32", "output_indptr", True))


def collect_compressed_float32_t_int32_t_int32_t(input_data, input_indices, input_indptr, output_data, output_indices,
                                                 output_indptr):
    """relayout.cpp:73-108; `output_indptr` is the caller's ``indptr[1:]`` holding the band starts."""
    if len(output_data) != len(input_data) or len(output_indices) != len(input_indices):
        raise ValueError("output arrays must match the input sizes")
    call("mcb_host_collect_compressed_f32_i32_i32", host_ptr(input_data, "float32", "input_data"),
         host_ptr(input_indices, "int32", "input_indices"), host_ptr(input_indptr, "int32", "input_indptr"),
         ("i64", len(input_indptr) - 1), host_ptr(output_data, "float32", "output_data", True),
         host_ptr(output_indices, "int32", "output_indices", True), host_ptr(output_indptr, "int32", "output_indptr", True),
         ("i64", len(output_indptr)))


def sort_compressed_indices_float32_t_int32_t_int32_t(data, indices, indptr, elements_count):
    """relayout.cpp:152-167: in place."""
    del elements_count  # only used by the reference for an assertion
    call("mcb_host_sort_compressed_indices_f32_i32_i32", host_ptr(data, "float32", "data", True),
         host_ptr(indices, "int32", "indices", True), host_ptr(indptr, "int32", "indptr"), ("i64", len(indptr) - 1))


# @@APPEND@@
