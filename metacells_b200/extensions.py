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


# @@APPEND@@
