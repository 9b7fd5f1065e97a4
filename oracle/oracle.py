"""
oracle.py -- TEST INFRASTRUCTURE ONLY (see the header of mc_oracle.c).

ctypes front-end of ``oracle/libmc_oracle.so`` (the plain-C restatement of the reference's native
functions) plus a numpy/scipy restatement of the Python glue between the native calls on the
per-pile k-NN path.  Every function cites the reference file:line it follows (paths relative to
the tanaylab/metacells v0.9.5 tree).  Nothing under ``metacells_b200/`` may import this module.
"""
from __future__ import annotations

import ctypes
import math
import os
import subprocess
from typing import Optional, Tuple

import numpy as np
import scipy.sparse as sp

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libmc_oracle.so")
_lib: Optional[ctypes.CDLL] = None


def build(force: bool = False) -> str:
    """Compile mc_oracle.c + mc_partitions.c + mc_seeds.c with gcc (seconds).  Building the checker is not using it."""
    sources = [os.path.join(_HERE, name) for name in ("mc_oracle.c", "mc_partitions.c", "mc_seeds.c")]
    newest = max(os.path.getmtime(src) for src in sources)
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < newest:
        cc = "/usr/bin/gcc" if os.path.exists("/usr/bin/gcc") else "gcc"
        subprocess.check_call([cc, "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-o", _LIB_PATH, *sources, "-lm"])
    return _LIB_PATH


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(build())
    return _lib


def _p(array: np.ndarray, dtype: str) -> ctypes.c_void_p:
    assert array.dtype == np.dtype(dtype), (array.dtype, dtype)
    assert array.flags.c_contiguous
    return ctypes.c_void_p(array.ctypes.data)


_i64 = ctypes.c_int64
_u64 = ctypes.c_uint64
_i32 = ctypes.c_int32


def _check(code: int) -> None:
    if code != 0:
        raise MemoryError("mc_oracle: allocation failed")


# ---------------------------------------------------------------------------------------------
# native restatements
# ---------------------------------------------------------------------------------------------


def downsample_compressed(data: np.ndarray, indptr: np.ndarray, samples: np.ndarray, seed: int) -> np.ndarray:
    """downsample.cpp:171-192; returns the downsampled ``data`` (same sparsity structure)."""
    out = np.zeros_like(data)
    _check(lib().mco_downsample_compressed_f32_i32_f32(
        _p(data, "float32"), _p(indptr, "int32"), _p(out, "float32"), _p(samples, "int32"), _i64(len(indptr) - 1), _u64(seed)))
    return out


def downsample_dense(matrix: np.ndarray, samples: np.ndarray, seed: int) -> np.ndarray:
    """downsample.cpp:131-151."""
    out = np.zeros_like(matrix)
    _check(lib().mco_downsample_dense_f32_f32(
        _p(matrix, "float32"), _p(out, "float32"), _p(samples, "int32"), _i64(matrix.shape[0]), _i64(matrix.shape[1]), _u64(seed)))
    return out


def downsample_array(vector: np.ndarray, samples: int, seed: int) -> np.ndarray:
    """downsample.cpp:116-128."""
    out = np.zeros_like(vector)
    _check(lib().mco_downsample_array_f32_f32(_p(vector, "float32"), _p(out, "float32"), _i64(len(vector)), _i32(samples), _u64(seed)))
    return out


def correlate_dense(dense: np.ndarray) -> np.ndarray:
    """correlate.cpp:350-431."""
    out = np.empty((dense.shape[0], dense.shape[0]), dtype="float32")
    _check(lib().mco_correlate_dense_f32(_p(dense, "float32"), _i64(dense.shape[0]), _i64(dense.shape[1]), _i64(dense.shape[1]), _p(out, "float32")))
    return out


def cross_correlate_dense(first: np.ndarray, second: np.ndarray) -> np.ndarray:
    """correlate.cpp:433-494."""
    out = np.empty((first.shape[0], second.shape[0]), dtype="float32")
    _check(lib().mco_cross_correlate_dense_f32(
        _p(first, "float32"), _i64(first.shape[0]), _p(second, "float32"), _i64(second.shape[0]), _i64(first.shape[1]), _p(out, "float32")))
    return out


def query_atlas_candidates(query: np.ndarray, atlas: np.ndarray, count: int) -> Tuple[np.ndarray, np.ndarray]:
    """tools/project.py:270 + :310: the cross correlation and, per query row, the head of ``np.argsort(-row)`` (made
    stable: numpy's default quicksort leaves the order of equal correlations unspecified)."""
    correlations = cross_correlate_dense(query, atlas)
    order = np.argsort(-correlations, axis=1, kind="stable")[:, :count]
    return order.astype("int32"), np.take_along_axis(correlations, order, axis=1)


def pairs_correlate_dense(first: np.ndarray, second: np.ndarray) -> np.ndarray:
    """correlate.cpp:496-527."""
    out = np.empty(first.shape[0], dtype="float32")
    _check(lib().mco_pairs_correlate_dense_f32(_p(first, "float32"), _p(second, "float32"), _i64(first.shape[0]), _i64(first.shape[1]), _p(out, "float32")))
    return out


def rank_matrix(matrix: np.ndarray, ascending: bool) -> np.ndarray:
    """rank.cpp:79-87 (returns a ranked copy)."""
    out = np.array(matrix, dtype="float32", order="C", copy=True)
    _check(lib().mco_rank_matrix_f32(_p(out, "float32"), _i64(out.shape[0]), _i64(out.shape[1]), ctypes.c_int(int(ascending))))
    return out


def rank_rows(matrix: np.ndarray, rank: int) -> np.ndarray:
    """rank.cpp:24-36."""
    out = np.empty(matrix.shape[0], dtype="float32")
    _check(lib().mco_rank_rows_f32(_p(matrix, "float32"), _i64(matrix.shape[0]), _i64(matrix.shape[1]), _p(out, "float32"), _i64(rank)))
    return out


def collect_top(degree: int, similarity: np.ndarray, ranks: bool) -> Tuple[np.ndarray, np.ndarray]:
    """top_per.cpp:67-92; returns (indices int32[N*degree], data float32[N*degree])."""
    rows = similarity.shape[0]
    indices = np.empty(rows * degree, dtype="int32")
    data = np.empty(rows * degree, dtype="float32")
    _check(lib().mco_collect_top_f32(_i64(degree), _p(similarity, "float32"), _i64(rows), _i64(similarity.shape[1]),
                                     _p(indices, "int32"), _p(data, "float32"), ctypes.c_int(int(ranks))))
    return indices, data


def collect_pruned(degree: int, data: np.ndarray, indices: np.ndarray, indptr: np.ndarray) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
    """prune_per.cpp:56-105; returns (data, indices, indptr) trimmed to indptr[-1]."""
    bands = len(indptr) - 1
    out_data = np.empty(max(bands * degree, 1), dtype="float32")
    out_indices = np.empty(max(bands * degree, 1), dtype="int32")
    out_indptr = np.empty(bands + 1, dtype="int32")
    _check(lib().mco_collect_pruned_f32_i32_i32(_i64(degree), _p(data, "float32"), _p(indices, "int32"), _p(indptr, "int32"),
                                                _i64(bands), _p(out_data, "float32"), _p(out_indices, "int32"), _p(out_indptr, "int32")))
    nnz = int(out_indptr[-1])
    return out_data[:nnz], out_indices[:nnz], out_indptr


def relayout(data: np.ndarray, indices: np.ndarray, indptr: np.ndarray, bands_out: int) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
    """relayout.cpp:73-108 + :152-167 as sequenced by computation.py:251-303."""
    out_data = np.empty_like(data)
    out_indices = np.empty_like(indices)
    out_indptr = np.empty(bands_out + 1, dtype="int32")
    _check(lib().mco_relayout_f32_i32_i32(_p(data, "float32"), _p(indices, "int32"), _p(indptr, "int32"), _i64(len(indptr) - 1),
                                          _i64(bands_out), _p(out_data, "float32"), _p(out_indices, "int32"), _p(out_indptr, "int32")))
    return out_data, out_indices, out_indptr


# ---------------------------------------------------------------------------------------------
# numpy / scipy glue between the native calls
# ---------------------------------------------------------------------------------------------


def downsample_samples(total_per_cell: np.ndarray, min_samples: int = 750, min_quantile: float = 0.05, max_quantile: float = 0.5) -> int:
    """tools/downsample.py:84-93 (defaults parameters.py: downsample_min_samples etc.)."""
    return int(round(min(max(min_samples, np.quantile(total_per_cell, min_quantile)), np.quantile(total_per_cell, max_quantile))))


def log_fraction(dense: np.ndarray, regularization: float = 1 / 7) -> np.ndarray:
    """pipeline/direct.py:242-249 + utilities/computation.py:679-741 (base 2, normalization 0)."""
    data = np.array(dense, dtype="float32", copy=True)
    data += regularization
    np.log2(data, out=data)
    return data


def knn_k(n_obs: int, cell_umis: np.ndarray, target_metacell_size: int = 48, target_metacell_umis: int = 160000,
          knn_k_size_factor: float = 2.0, knn_k_umis_quantile: float = 0.1, min_knn_k: int = 30) -> int:
    """pipeline/direct.py:251-267."""
    by_size = int(round(n_obs / target_metacell_size))
    by_median = int(round(knn_k_size_factor * target_metacell_umis / float(np.median(cell_umis))))
    by_quantile = int(round(target_metacell_umis / np.quantile(cell_umis, knn_k_umis_quantile)))
    return max(by_size, by_median, by_quantile, min_knn_k or 0)


def similarity(dense: np.ndarray, method: str = "abs_pearson") -> np.ndarray:
    """utilities/computation.py:454-467 (reproducible corrcoef) + tools/similarity.py:213-215."""
    result = correlate_dense(np.ascontiguousarray(dense, dtype="float32"))
    result[np.isnan(result)] = 0.0
    np.fill_diagonal(result, 1.0)
    if method == "abs_pearson":
        np.absolute(result, out=result)
    return result


def rank_outgoing(sim: np.ndarray) -> np.ndarray:
    """tools/knn_graph.py:285-297."""
    work = np.copy(sim)
    np.fill_diagonal(work, np.min(work) - 1)
    ranks = rank_matrix(work, ascending=False)
    assert np.all(np.diagonal(ranks) == sim.shape[0])
    return ranks


def balance_ranks(outgoing_ranks: np.ndarray, k: int, balanced_ranks_factor: float) -> sp.csr_matrix:
    """tools/knn_graph.py:300-340.  The float32 operation order (multiply, sqrt, the 2**21 offset
    round trip) is kept because it rounds."""
    size = outgoing_ranks.shape[0]
    dense = np.array(outgoing_ranks, dtype="float32", copy=True)
    dense *= outgoing_ranks.transpose()
    np.sqrt(dense, out=dense)
    max_rank = k * balanced_ranks_factor
    dense *= -1
    dense += 2**21
    max_index_of_each = np.asarray(dense.argmax(axis=1)).reshape(-1)
    dense += max_rank + 1 - 2**21
    rows = np.arange(size)
    preserved = dense[rows, max_index_of_each]
    preserved[preserved < 1] = 1
    dense[dense < 0] = 0
    np.fill_diagonal(dense, 0)
    dense[rows, max_index_of_each] = preserved
    return sp.csr_matrix(dense)


def prune_per(compressed, top: int):
    """utilities/computation.py:1356-1393."""
    is_csr = compressed.format == "csr"
    data, indices, indptr = collect_pruned(top, compressed.data, compressed.indices, compressed.indptr)
    constructor = sp.csr_matrix if is_csr else sp.csc_matrix
    return constructor((data, indices, indptr), shape=compressed.shape)


def csc_to_csr(csc: sp.csc_matrix) -> sp.csr_matrix:
    """utilities/computation.py:251-303 (_relayout_compressed)."""
    data, indices, indptr = relayout(csc.data, csc.indices, csc.indptr, csc.shape[0])
    return sp.csr_matrix((data, indices, indptr), shape=csc.shape)


def csr_to_csc(csr: sp.csr_matrix) -> sp.csc_matrix:
    data, indices, indptr = relayout(csr.data, csr.indices, csr.indptr, csr.shape[1])
    return sp.csc_matrix((data, indices, indptr), shape=csr.shape)


def prune_ranks(balanced: sp.csr_matrix, k: int, incoming_degree_factor: float, outgoing_degree_factor: float) -> sp.csr_matrix:
    """tools/knn_graph.py:343-392."""
    size = balanced.shape[0]
    incoming_degree = min(int(round(k * incoming_degree_factor)), size - 1)
    outgoing_degree = min(int(round(k * outgoing_degree_factor)), size - 1)
    rows = np.arange(size)
    max_index_of_each = np.asarray(balanced.argmax(axis=1)).reshape(-1)
    preserved_values = np.asarray(balanced[rows, max_index_of_each]).reshape(-1)
    assert np.min(preserved_values) > 0
    preserved = sp.coo_matrix((preserved_values, (rows, max_index_of_each)), shape=balanced.shape)
    # knn_graph.py:371: symmetric=True relayout is a reinterpretation of the same three arrays.
    pruned = sp.csc_matrix((balanced.data, balanced.indices, balanced.indptr), shape=balanced.shape)
    pruned = prune_per(pruned, incoming_degree)
    pruned = csc_to_csr(pruned)
    pruned = prune_per(pruned, outgoing_degree)
    pruned = pruned.maximum(preserved)
    pruned = pruned.maximum(preserved.transpose())
    pruned = sp.csr_matrix(pruned)
    pruned.sort_indices()
    return pruned


def weigh_edges(pruned: sp.csr_matrix) -> sp.csr_matrix:
    """tools/knn_graph.py:395-407."""
    totals = np.asarray(pruned.sum(axis=1)).reshape(-1).astype("float32")
    scale = np.reciprocal(totals)
    weights = sp.csr_matrix(pruned.multiply(scale[:, None]))
    weights.sort_indices()
    return weights


def knn_graph(sim: np.ndarray, k: int, balanced_ranks_factor: float = float(np.sqrt(10)), incoming_degree_factor: float = 3.0,
              outgoing_degree_factor: float = 1.0, min_outgoing_degree: int = 2):
    """tools/knn_graph.py:251-275 (the k-growth loop).  Returns (outgoing_weights, pruned_ranks, balanced_ranks, k)."""
    while True:
        ranks = rank_outgoing(sim)
        balanced = balance_ranks(ranks, k, balanced_ranks_factor)
        pruned = prune_ranks(balanced, k, incoming_degree_factor, outgoing_degree_factor)
        if np.min(np.diff(pruned.indptr)) >= min_outgoing_degree:
            break
        k += max(1, int(k / 10))
    return weigh_edges(pruned), pruned, balanced, k


# ---------------------------------------------------------------------------------------------
# partitions.cpp (mc_partitions.c)
# ---------------------------------------------------------------------------------------------


class ReferenceAssert(RuntimeError):
    """A FastAssert of the reference failed (the reference would print it and exit(1))."""


def _partition_args(outgoing: sp.csr_matrix, incoming: sp.csc_matrix):
    assert outgoing.shape[0] == outgoing.shape[1] == incoming.shape[0] == incoming.shape[1]
    return (_p(outgoing.data, "float32"), _p(outgoing.indices, "int32"), _p(outgoing.indptr, "int32"),
            _p(incoming.data, "float32"), _p(incoming.indices, "int32"), _p(incoming.indptr, "int32"), _i64(outgoing.shape[0]))


def optimize_partitions(outgoing: sp.csr_matrix, incoming: sp.csc_matrix, random_seed: int, mass_of_nodes: np.ndarray,
                        low_mass: float, target_mass: float, high_mass: float, low_size: float, target_size: float,
                        high_size: float, cooldown_pass: float, cooldown_node: float, partition_of_nodes: np.ndarray,
                        hot_partitions_mask: np.ndarray, cold_temperature: float, stats: Optional[np.ndarray] = None) -> float:
    """partitions.cpp:1234-1302; ``partition_of_nodes`` is modified in place, the score is returned.
    ``stats`` (int64[4], optional) receives passes / improve_node calls / moves / skipped visits."""
    bounds = np.array([low_mass, target_mass, high_mass, low_size, target_size, high_size], dtype="float64")
    score = ctypes.c_double()
    if stats is not None:
        stats[:] = 0
    function = lib().mco_optimize_partitions
    function.restype = ctypes.c_int
    status = function(*_partition_args(outgoing, incoming), ctypes.c_uint32(random_seed % (1 << 32)),
                      _p(mass_of_nodes, "float32"), _p(bounds, "float64"), ctypes.c_double(cooldown_pass),
                      ctypes.c_double(cooldown_node), _p(partition_of_nodes, "int32"), _p(hot_partitions_mask, "int8"),
                      ctypes.c_double(cold_temperature), ctypes.byref(score),
                      _p(stats, "int64") if stats is not None else ctypes.c_void_p())
    if status == 2:
        raise ReferenceAssert("optimize_partitions: a FastAssert of the reference fails on this input")
    _check(status)
    return score.value


def score_partitions(outgoing: sp.csr_matrix, incoming: sp.csc_matrix, mass_of_nodes: np.ndarray, low_mass: float,
                     target_mass: float, high_mass: float, low_size: float, target_size: float, high_size: float,
                     partition_of_nodes: np.ndarray, hot_partitions_mask: np.ndarray, with_orphans: bool = True) -> float:
    """partitions.cpp:1304-1338."""
    bounds = np.array([low_mass, target_mass, high_mass, low_size, target_size, high_size], dtype="float64")
    score = ctypes.c_double()
    function = lib().mco_score_partitions
    function.restype = ctypes.c_int
    status = function(*_partition_args(outgoing, incoming), _p(mass_of_nodes, "float32"), _p(bounds, "float64"),
                      _p(partition_of_nodes, "int32"), _p(hot_partitions_mask, "int8"), ctypes.c_int(int(with_orphans)),
                      ctypes.byref(score))
    if status == 2:
        raise ReferenceAssert("score_partitions: a FastAssert of the reference fails on this input")
    _check(status)
    return score.value


def choose_seeds(outgoing: sp.csr_matrix, incoming: sp.csc_matrix, random_seed: int, max_seeds_count: int,
                 min_seed_size_quantile: float, max_seed_size_quantile: float, seed_of_nodes: np.ndarray) -> int:
    """choose_seeds.cpp:264-372; ``seed_of_nodes`` (int32, -1 = no seed yet) is filled in place, the number of
    seeds is returned."""
    count = ctypes.c_int64()
    function = lib().mco_choose_seeds
    function.restype = ctypes.c_int
    status = function(*_partition_args(outgoing, incoming), _u64(random_seed), _i64(max_seeds_count),
                      ctypes.c_float(min_seed_size_quantile), ctypes.c_float(max_seed_size_quantile),
                      _p(seed_of_nodes, "int32"), ctypes.byref(count))
    if status == 2:
        raise ReferenceAssert("choose_seeds: a FastAssert of the reference fails on this input")
    _check(status)
    return count.value


def std_nth_element(keys: np.ndarray, first: int, nth: int, last: int, positions: Optional[np.ndarray] = None) -> np.ndarray:
    """``std::nth_element`` (libstdc++ 13) over positions compared by ``keys[position]``; returns the positions."""
    keys = np.ascontiguousarray(keys, dtype="int64")
    positions = np.arange(len(keys), dtype="int64") if positions is None else positions
    lib().mco_nth_element_positions(_p(positions, "int64"), _p(keys, "int64"), _i64(first), _i64(nth), _i64(last))
    return positions


def std_sort(keys: np.ndarray) -> np.ndarray:
    """``std::sort`` (libstdc++ 13) of ``iota(len(keys))`` compared by ``keys[position]``; returns the positions."""
    keys = np.ascontiguousarray(keys, dtype="int64")
    positions = np.arange(len(keys), dtype="int64")
    lib().mco_sort_positions(_p(positions, "int64"), _p(keys, "int64"), _i64(len(keys)))
    return positions


def glibc_log(values: np.ndarray) -> np.ndarray:
    """libm's double ``log`` (what extensions.h:135-138 calls), element by element."""
    values = np.ascontiguousarray(values, dtype="float64")
    out = np.empty_like(values)
    lib().mco_log_array(_p(values, "float64"), _p(out, "float64"), _i64(values.size))
    return out


def std_shuffle(count: int, seed: int, rounds: int = 1) -> np.ndarray:
    """``std::shuffle`` of ``iota(count)`` with ``std::minstd_rand(seed)``, applied ``rounds`` times (libstdc++ 13)."""
    indices = np.arange(count, dtype="int64")
    lib().mco_shuffle(_p(indices, "int64"), _i64(count), ctypes.c_uint32(seed), _i64(rounds))
    return indices


# ---------------------------------------------------------------------------------------------
# the compiled reference (oracle/_ref), when present
# ---------------------------------------------------------------------------------------------


def load_ref():
    """Import oracle/_ref/extensions*.so (the unmodified reference C++), or return None."""
    import glob
    import importlib.util

    found = glob.glob(os.path.join(_HERE, "_ref", "extensions*.so"))
    if not found:
        return None
    spec = importlib.util.spec_from_file_location("extensions", found[0])
    module = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(module)
    return module


# ---------------------------------------------------------------------------------------------
# the whole per-pile path on the CPU (bench.py's cpu_baseline / --impl reference leg)
# ---------------------------------------------------------------------------------------------


def cpu_pile_knn_graph(umis: sp.csr_matrix, features, k: int, seed: int, ref=None, timings: Optional[dict] = None):
    """downsample -> feature genes -> log-fraction -> abs_pearson similarity -> balanced k-NN graph,
    sequenced as pipeline/direct.py:214-283 does.  With ``ref`` (oracle/_ref, the unmodified reference
    natives, multi-threaded) every native call goes to the reference; otherwise to the C port of
    this directory (single thread).  Returns the outgoing weights CSR matrix."""
    import time

    def lap(name, start):
        if timings is not None:
            timings[name] = timings.get(name, 0.0) + time.perf_counter() - start

    start = time.perf_counter()
    totals = np.asarray(umis.sum(axis=1)).reshape(-1)
    samples = np.full(umis.shape[0], downsample_samples(totals), dtype="int32")
    if ref is not None:
        down_data = np.empty_like(umis.data)
        ref.downsample_compressed_float32_t_int32_t_float32_t(umis.data, umis.indptr, down_data, samples, seed)
    else:
        down_data = downsample_compressed(umis.data, umis.indptr, samples, seed)
    down = sp.csr_matrix((down_data, umis.indices, umis.indptr), shape=umis.shape)
    lap("downsample", start)

    start = time.perf_counter()
    if isinstance(features, str):  # "select": extract_selected_data of pipeline/select.py:145-220
        assert features == "select"
        stored = sp.csr_matrix(down, copy=True)
        stored.eliminate_zeros()  # computation.py:892-899
        genes = np.flatnonzero(select_feature_genes(stored, ref=ref))
        del stored
        lap("select_genes", start)
        start = time.perf_counter()
    else:  # the stand-in of SURVEY.md 8d: the genes of the largest downsampled totals
        gene_totals = np.asarray(down.sum(axis=0)).reshape(-1)
        genes = np.sort(np.argsort(-gene_totals, kind="stable")[:features])
    if timings is not None:
        timings["feature_genes_count"] = len(genes)
    dense = log_fraction(down[:, genes].toarray())
    lap("log_fraction", start)

    start = time.perf_counter()
    if ref is not None:
        sim = np.empty((dense.shape[0], dense.shape[0]), dtype="float32")
        ref.correlate_dense_float32_t(dense, sim)
    else:
        sim = correlate_dense(dense)
    sim[np.isnan(sim)] = 0.0
    np.fill_diagonal(sim, 1.0)
    np.absolute(sim, out=sim)
    lap("correlate", start)

    def ref_prune(compressed, top):
        bands = compressed.shape[0]
        out_data = np.empty(bands * top, dtype="float32")
        out_indices = np.empty(bands * top, dtype="int32")
        out_indptr = np.empty(bands + 1, dtype="int32")
        ref.collect_pruned_float32_t_int32_t_int32_t(top, compressed.data, compressed.indices, compressed.indptr, out_data, out_indices, out_indptr)
        nnz = out_indptr[-1]
        constructor = sp.csr_matrix if compressed.format == "csr" else sp.csc_matrix
        return constructor((out_data[:nnz], out_indices[:nnz], out_indptr), shape=compressed.shape)

    def ref_csc_to_csr(csc):
        size = csc.shape[0]
        counts = np.bincount(csc.indices, minlength=size)
        indptr = np.zeros(size + 1, dtype="int32")
        np.cumsum(counts[:-1], out=indptr[2:])
        data = np.empty_like(csc.data)
        indices = np.empty_like(csc.indices)
        ref.collect_compressed_float32_t_int32_t_int32_t(csc.data, csc.indices, csc.indptr, data, indices, indptr[1:])
        ref.sort_compressed_indices_float32_t_int32_t_int32_t(data, indices, indptr, size)
        return sp.csr_matrix((data, indices, indptr), shape=csc.shape)

    size = sim.shape[0]
    while True:
        start = time.perf_counter()
        if ref is not None:
            ranks = np.copy(sim)
            np.fill_diagonal(ranks, np.min(ranks) - 1)
            ref.rank_matrix_float32_t(ranks, False)
        else:
            ranks = rank_outgoing(sim)
        lap("rank_matrix", start)
        start = time.perf_counter()
        balanced = balance_ranks(ranks, k, float(np.sqrt(10)))
        lap("balance_ranks", start)
        start = time.perf_counter()
        if ref is not None:
            incoming = min(int(round(k * 3.0)), size - 1)
            outgoing = min(int(round(k * 1.0)), size - 1)
            rows = np.arange(size)
            best = np.asarray(balanced.argmax(axis=1)).reshape(-1)
            preserved = sp.coo_matrix((np.asarray(balanced[rows, best]).reshape(-1), (rows, best)), shape=balanced.shape)
            pruned = sp.csc_matrix((balanced.data, balanced.indices, balanced.indptr), shape=balanced.shape)
            pruned = ref_prune(pruned, incoming)
            pruned = ref_csc_to_csr(pruned)
            pruned = ref_prune(pruned, outgoing)
            pruned = pruned.maximum(preserved).maximum(preserved.transpose())
            pruned = sp.csr_matrix(pruned)
            pruned.sort_indices()
        else:
            pruned = prune_ranks(balanced, k, 3.0, 1.0)
        lap("prune_ranks", start)
        if np.min(np.diff(pruned.indptr)) >= 2:
            break
        k += max(1, int(k / 10))
    start = time.perf_counter()
    weights = weigh_edges(pruned)
    lap("weigh_edges", start)
    return weights


# ---------------------------------------------------------------------------------------------
# feature-gene selection (pipeline/select.py:145-220 over tools/high.py and utilities/computation.py)
# ---------------------------------------------------------------------------------------------


def high_topN_genes(downsampled: sp.spmatrix, topN: int, min_gene_topN: int, ref=None) -> np.ndarray:
    """tools/high.py:121-124: ``rank_per(column_major, rank=max(n_obs - topN - 1, 1)) >= min_gene_topN``
    (computation.py:1285-1293 -> rank_rows on the dense [genes, cells] view)."""
    dense = np.ascontiguousarray(downsampled.toarray().transpose(), dtype="float32")
    rank = max(downsampled.shape[0] - topN - 1, 1)
    if ref is not None:
        values = np.empty(dense.shape[0], dtype="float32")
        ref.rank_rows_float32_t(dense, values, rank)
    else:
        values = rank_rows(dense, rank)
    return values >= min_gene_topN


def high_total_genes(downsampled: sp.spmatrix, min_gene_total: float) -> np.ndarray:
    """tools/high.py:29-79: ``sum_per(column) >= min_gene_total``."""
    return np.asarray(sp.csc_matrix(downsampled).sum(axis=0)).reshape(-1) >= min_gene_total


def _variance_per_column(matrix: sp.csc_matrix) -> np.ndarray:
    """utilities/computation.py:1526-1546 with sum_per / sum_squared_per (:1224-1268) on a CSC matrix."""
    sums = np.asarray(matrix.sum(axis=0)).reshape(-1)
    sums_squared = np.asarray(matrix.multiply(matrix).sum(axis=0)).reshape(-1)
    size = matrix.shape[0]
    result = np.square(sums).astype(float)
    result /= -size
    result += sums_squared
    result /= size
    result[result < 0] = 0
    return result


def relative_variance_per_column(matrix: sp.spmatrix, window_size: int = 100) -> np.ndarray:
    """utilities/computation.py:1549-1595 (normalized_variance_per, relative_variance_per) and :1836-1897
    (sliding_window_function, ``median``, ordered by the mean)."""
    import pandas as pd

    matrix = sp.csc_matrix(matrix)
    variance = _variance_per_column(matrix)
    mean = np.asarray(matrix.sum(axis=0)).reshape(-1) / matrix.shape[0]
    zeros_mask = mean == 0
    normalized = np.reciprocal(mean, where=~zeros_mask)
    normalized[zeros_mask] = 0
    normalized *= variance
    normalized[zeros_mask] = 1.0
    with np.errstate(divide="ignore"):
        np.log2(normalized, out=normalized)
    if window_size % 2 == 0:
        window_size += 1
    half = (window_size - 1) // 2
    size = len(normalized)
    if window_size >= size:
        return normalized - np.full(size, np.median(normalized))
    order = np.argsort(mean)
    reverse = np.argsort(order)
    extended = np.concatenate([order[np.arange(window_size - half, window_size)], order,
                               order[np.arange(size - window_size, size - window_size + half)]])
    rolled = pd.Series(normalized[extended]).rolling(window_size).median().values[window_size - 1:]
    assert rolled.size == size
    return normalized - rolled[reverse]


def select_feature_genes(downsampled: sp.spmatrix, *, min_gene_relative_variance: Optional[float] = 0.1, min_genes: int = 100,
                         min_gene_total: Optional[int] = 50, min_gene_top3: Optional[int] = 4,
                         lateral_genes_mask: Optional[np.ndarray] = None, ref=None) -> np.ndarray:
    """pipeline/select.py:145-220 (no ``select_gene`` override): the boolean ``selected_gene`` mask."""
    genes = downsampled.shape[1]
    masks = []
    if min_gene_top3 is not None:
        masks.append(high_topN_genes(downsampled, 3, min_gene_top3, ref=ref))
    if min_gene_total is not None:
        masks.append(high_total_genes(downsampled, min_gene_total))
    relative = None
    if min_gene_relative_variance is not None:
        relative = relative_variance_per_column(downsampled)
        masks.append(relative >= min_gene_relative_variance)
    additional = [] if lateral_genes_mask is None else [~np.asarray(lateral_genes_mask, dtype=bool)]

    def combine(some):
        combined = np.ones(genes, dtype=bool)
        for mask in some:
            combined &= mask
        return combined

    candidates = combine(masks + additional)
    if np.sum(candidates) >= min_genes:
        return candidates
    if min_gene_relative_variance is None:
        raise ValueError(f"Failed to select {min_genes} genes")
    valid = combine(masks[:-1] + additional)
    is_valid_set = False
    if np.sum(valid) <= min_genes:
        masks.pop()
    else:
        while True:
            high = min_gene_relative_variance
            min_gene_relative_variance -= 1 / 8
            masks[-1] = relative >= min_gene_relative_variance
            valid = combine(masks + additional)
            if np.sum(valid) >= min_genes:
                is_valid_set = True
                break
        for _ in range(4):
            mid = (high + min_gene_relative_variance) / 2.0
            masks[-1] = relative >= mid
            mid_mask = combine(masks + additional)
            if np.sum(mid_mask) >= min_genes:
                min_gene_relative_variance = mid
                valid = mid_mask
                is_valid_set = True
            else:
                high = mid
                is_valid_set = False
        # select.py:208-209: the last mask written (the mid one) stays unless the valid set is put back
        if not is_valid_set:
            masks[-1] = valid
    return combine(masks + additional)


# ---- the candidate-metacells loop (tools/candidates.py:32-289) ----------------------------------------------------------
#
# Host control flow of the reference restated around this file's choose_seeds / optimize_partitions (the C restatements
# pinned against oracle/_ref).  ``ut.min_cut`` hands the community's edges to igraph's ``Graph.mincut``, which is NOT in
# this image: ``stoer_wagner_min_cut`` below is the published algorithm igraph implements for undirected graphs, with this
# file's own tie rules, so the CHOICE among several cuts of the same minimal weight is unpinned against igraph (the weight
# of the cut, and so the strength test, is not affected when the minimal cut is unique).


def stoer_wagner_min_cut(weights: np.ndarray) -> np.ndarray:
    """The side (``True`` = ``cut.partition[1]``) of every vertex in a minimal cut of the undirected graph with the
    symmetric dense ``weights`` (Stoer & Wagner 1997), in float64.  Every phase grows a set from the lowest live vertex
    by the most tightly connected one (ties: the lowest index); the first phase with the least cut-of-the-phase wins and
    the vertices merged into its last vertex are ``cut.partition[0]``."""
    count = len(weights)
    merged = np.array(weights, dtype="float64")
    root = np.arange(count)
    live = np.ones(count, dtype=bool)
    best, side = np.inf, np.zeros(count, dtype=bool)
    for phase in range(count - 1):
        key = np.zeros(count)
        free = live.copy()
        last = previous = -1
        last_key = 0.0
        for _ in range(count - phase):
            chosen = int(np.argmax(np.where(free, key, -1.0)))
            previous, last, last_key = last, chosen, key[chosen]
            free[chosen] = False
            key[free] += merged[chosen, free]
        if last_key < best:
            best, side = last_key, root != last
        others = live.copy()
        others[[last, previous]] = False
        merged[previous, others] += merged[last, others]
        merged[others, previous] = merged[previous, others]
        live[last] = False
        root[root == last] = previous
    return side


def min_cut_strength(weights: np.ndarray, is_second: np.ndarray) -> Optional[float]:
    """The strength ``ut.min_cut`` returns beside the cut (utilities/computation.py:2408-2449).  The three totals are
    ``total += weight`` of numpy float32 scalars, which under NumPy >= 2 stays float32: sequential float32 sums, here as
    the last entry of a float32 ``cumsum``, over the edges in the order :2389-2403 lists them."""
    count = len(weights)
    sources, targets = np.nonzero(np.tril(weights, -1) > 0)  # row-major: source ascending, target ascending below it
    edge_weights = weights[sources, targets].astype("float32")
    across = is_second[sources] != is_second[targets]
    inside_second = is_second[sources] & is_second[targets]
    inside_first = ~is_second[sources] & ~is_second[targets]

    def total(mask):
        return np.cumsum(edge_weights[mask], dtype="float32")[-1] if mask.any() else 0.0

    cut_total, first_total, second_total = total(across), total(inside_first), total(inside_second)
    second_size = int(is_second.sum())
    first_size = count - second_size
    if cut_total == 0:
        return 0.0
    if not (first_total > 0 and second_total > 0):
        return None
    cut_mean = cut_total / (first_size * second_size)
    first_mean = first_total / ((first_size * (first_size - 1)) / 2)
    second_mean = second_total / ((second_size * (second_size - 1)) / 2)
    return cut_mean / math.sqrt(first_mean * second_mean)


def _candidate_seeds_count(total_umis, total_size, low_umis, high_umis, low_size, high_size) -> int:
    """tools/candidates.py:293-321."""
    by_umis = (int(math.ceil(total_umis / high_umis)), int(math.ceil(total_umis / low_umis)))
    by_size = (int(math.ceil(total_size / high_size)), int(math.ceil(total_size / low_size)))
    if by_size[1] < by_umis[0]:
        return int(round((by_size[1] + by_umis[0]) / 2))
    if by_umis[1] < by_size[0]:
        return int(round((by_umis[1] + by_size[0]) / 2))
    return int(round((max(by_size[0], by_umis[0]) + min(by_size[1], by_umis[1])) / 2))


def compute_candidate_metacells(outgoing: sp.csr_matrix, cell_umis: np.ndarray, *, random_seed: int,
                                target_metacell_size: int = 48, min_metacell_size: int = 12, target_metacell_umis: int = 160000,
                                min_seed_size_quantile: float = 0.85, max_seed_size_quantile: float = 0.95,
                                cooldown_pass: float = 0.02, cooldown_node: float = 0.25, cooldown_phase: float = 0.75,
                                increase_phase: float = 1.01, min_split_size_factor: float = 2.0,
                                max_merge_size_factor: float = 0.5, max_split_min_cut_strength: float = 0.1,
                                min_cut_seed_cells: int = 7, must_complete_cover: bool = False,
                                trace: Optional[list] = None) -> np.ndarray:
    """tools/candidates.py:121-289 with its helpers (:324-648) inlined as closures; the ``candidate`` of every cell."""
    incoming = csr_to_csc(outgoing)
    cells = outgoing.shape[0]
    assert cell_umis.dtype == np.float32
    lowest_size = min_metacell_size
    low_umis = int(math.floor(target_metacell_umis * max_merge_size_factor))  # :152-159
    high_umis = int(math.ceil(target_metacell_umis * min_split_size_factor)) - 1
    low_size = max(min_metacell_size, int(math.floor(target_metacell_size * max_merge_size_factor)))
    high_size = max(int(math.ceil(target_metacell_size * min_split_size_factor)) - 1, low_size + 1)
    known_whole = set()

    def seed(communities, max_seeds_count):  # :720-749
        count = choose_seeds(outgoing, incoming, random_seed, max_seeds_count, min_seed_size_quantile, max_seed_size_quantile,
                             communities)
        assert communities.min() == 0
        return count

    def optimize(communities, hot, cold_temperature, size_limit):  # :831-880
        hot_mask = np.zeros(max(communities.max() + 1, 2), dtype="int8")
        hot_mask[hot] = 1
        return optimize_partitions(outgoing, incoming, random_seed, cell_umis, low_umis, target_metacell_umis, high_umis,
                                   low_size, target_metacell_size, size_limit, cooldown_pass, cooldown_node, communities,
                                   hot_mask, cold_temperature)

    def cut_one(communities, index, new_index):  # :533-589
        mask = communities == index
        if mask.sum() < 2:
            return "unchanged"
        sliced = outgoing[mask, :][:, mask]
        weights = (sliced + sliced.T).toarray()
        is_second = stoer_wagner_min_cut(weights)
        strength = min_cut_strength(weights, is_second)
        if strength is None or strength > max_split_min_cut_strength:
            return "unchanged"
        where = np.where(mask)[0]
        sizes = (int((~is_second).sum()), int(is_second.sum()))
        small = 0 if sizes[0] < sizes[1] else 1
        if sizes[small] >= min_cut_seed_cells:
            communities[where[is_second]] = new_index
            return "split"
        if must_complete_cover and strength > 0:
            return "unchanged"
        communities[where[is_second == bool(small)]] = -1
        return "cut"

    def split_or_cut(communities, size_limit):  # :410-530
        count = communities.max() + 1
        new_index = count
        split = cut = 0
        hot = []
        for index in range(count):
            mask = communities == index
            size = np.sum(mask)
            umis = np.sum(cell_umis[mask])
            if not ((umis > high_umis and size >= 3 * low_size) or (size > size_limit and umis >= 3 * low_umis)):
                continue
            where = np.where(mask)[0]
            parts = max(2, min(int(round(umis / target_metacell_umis)), int(round(size / target_metacell_size))))
            while True:
                assignment = np.random.randint(parts, size=len(where))
                if all(np.any(assignment == part) for part in range(parts)):
                    break
            communities[where] = -1
            for part in range(parts):
                chosen = where[assignment == part]
                chosen = chosen[np.random.randint(len(chosen))]
                if part == 0:
                    hot.append(index)
                    communities[chosen] = index
                else:
                    hot.append(new_index)
                    communities[chosen] = new_index
                    new_index += 1
                    split += 1
        if split == 0:
            for index in range(count):
                key = tuple(np.where(communities == index)[0])
                if key in known_whole:
                    continue
                action = cut_one(communities, index, new_index)
                if action == "unchanged":
                    known_whole.add(key)
                    continue
                hot.append(index)
                if action == "split":
                    hot.append(new_index)
                    new_index += 1
                    split += 1
                else:
                    cut += 1
        return split, cut, hot

    def reduce(communities, hot, cold_temperature):  # :324-407
        np.random.seed(random_seed)
        size_limit = high_size
        while True:
            score = optimize(communities, hot, cold_temperature, size_limit)
            if trace is not None:
                trace.append(("optimize", score, communities.copy()))
            cold_temperature = cold_temperature * (1 - cooldown_phase)
            size_limit *= increase_phase
            split, cut, hot = split_or_cut(communities, size_limit)
            if trace is not None:
                trace.append(("split", split, cut, communities.copy()))
            if split + cut == 0:
                return cold_temperature, score
            max_seeds_count = communities.max() + 1
            if cut > 0:
                hot.append(max_seeds_count)
                max_seeds_count += 1
            seed(communities, max_seeds_count)

    def too_small(communities):  # :592-631
        small, nodes, umis_of_small = set(), 0, 0
        for index in range(communities.max() + 1):
            mask = communities == index
            size = np.sum(mask)
            umis = np.sum(cell_umis[mask])
            if size < lowest_size:
                pass
            elif size < low_size and 2 * umis < high_umis:
                pass
            elif umis < low_umis and 2 * size < high_size:
                pass
            else:
                continue
            small.add(index)
            nodes += size
            umis_of_small += umis
        return small, nodes, umis_of_small

    def cancel(communities, cancelled):  # :634-648
        kept = 0
        for index in range(communities.max() + 1):
            if index in cancelled:
                communities[communities == index] = -1
                continue
            if index > kept:
                communities[communities == index] = kept
            kept += 1
        return kept

    communities = np.full(cells, -1, dtype="int32")
    seeds_count = seed(communities, _candidate_seeds_count(sum(cell_umis), cells, low_umis, high_umis, low_size, high_size))
    if trace is not None:
        trace.append(("seeds", seeds_count, communities.copy()))
    np.random.seed(random_seed)
    cold_temperature = 1 - cooldown_pass
    old = (len(communities), 1e9)
    old_communities = communities
    hot = list(range(seeds_count))
    while True:
        cold_temperature, score = reduce(communities, hot, cold_temperature)
        small, small_nodes, small_umis = too_small(communities)
        if len(small) < 2:
            break
        if old <= (small_nodes, score):
            communities = old_communities
            break
        old, old_communities = (small_nodes, score), communities.copy()
        kept = cancel(communities, small)
        extra = min(len(small) - 1, _candidate_seeds_count(small_umis, small_nodes, low_umis, high_umis, low_size, high_size))
        seeds_count = seed(communities, kept + extra)
        if trace is not None:
            trace.append(("reseed", seeds_count, communities.copy()))
        hot = list(range(kept, seeds_count))
    assert np.all(communities >= 0)
    return communities


# ---- deviant cells, "gaps" policy (tools/deviants.py:241-356 over gaps.cpp) ----------------------------------------------


def gaps_deviant_cells(umis: sp.csr_matrix, candidates: np.ndarray, ref, *, noisy_genes_mask: Optional[np.ndarray] = None,
                       min_gene_fold_factor: float = 3.0, gap_skip_cells: int = 1, min_noisy_gene_fold_factor: float = 2.0,
                       max_cell_fraction: float = 0.25, max_gap_cells_count: int = 3, max_gap_cells_fraction: float = 0.1,
                       cells_regularization_quantile: float = 0.25, min_compare_umis: int = 8,
                       trace: Optional[list] = None) -> np.ndarray:
    """``find_deviant_cells(policy="gaps")``: the dense matrices as the reference makes them (numpy, on the host) and the
    reference's own ``compute_cell_gaps`` (``ref``: oracle/_ref, compiled from gaps.cpp).  ``gap_skip_cells`` is the
    public parameter (the native call gets one more, tools/deviants.py:100)."""
    gap_skip_cells += 1
    cells_count, genes_count = umis.shape
    if cells_count <= max_gap_cells_count:
        return np.zeros(cells_count, dtype="bool")
    min_gap_per_gene = np.full(genes_count, min_gene_fold_factor, dtype="float32")
    if noisy_genes_mask is not None:
        min_gap_per_gene[noisy_genes_mask] += min_noisy_gene_fold_factor
    candidate_per_cell = np.asarray(candidates).astype("int32")
    candidates_count = np.max(candidate_per_cell) + 1
    umis_per_gene_per_cell = np.ascontiguousarray(umis.toarray(), dtype="float32")
    regularization_per_cell = np.empty(cells_count, dtype="float32")
    max_gap_per_cell = np.empty(cells_count, dtype="float32")
    new_deviant_per_cell = np.empty(cells_count, dtype="bool")
    deviant_per_cell = np.empty(cells_count, dtype="bool")
    total_umis_per_cell = np.asarray(umis.sum(axis=1)).reshape(-1).copy()
    assert total_umis_per_cell.dtype == np.float32
    fraction_per_gene_per_cell = umis_per_gene_per_cell / total_umis_per_cell[:, np.newaxis]
    for candidate_index in range(candidates_count):
        mask_per_cell = candidate_per_cell == candidate_index
        totals = total_umis_per_cell[mask_per_cell]
        regularization = max(1.0, np.quantile(totals, cells_regularization_quantile))
        regularization_per_cell[mask_per_cell] = 1.0 / np.maximum(totals, regularization)
    log_fraction_per_gene_per_cell = np.log2(fraction_per_gene_per_cell + regularization_per_cell[:, np.newaxis])
    assert log_fraction_per_gene_per_cell.dtype == np.float32
    while True:
        new_deviant_per_cell[:] = True
        deviant_per_cell[:] = False
        deviants_fraction = 0.0
        while deviants_fraction <= max_cell_fraction:
            max_gap_per_cell[:] = 0.0
            ref.compute_cell_gaps(umis_per_gene_per_cell, fraction_per_gene_per_cell, log_fraction_per_gene_per_cell,
                                  candidate_per_cell, deviant_per_cell, new_deviant_per_cell, min_gap_per_gene, candidates_count,
                                  gap_skip_cells, max_gap_cells_count, max_gap_cells_fraction, float(min_compare_umis),
                                  max_gap_per_cell)
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


# ---- dissolve_metacells (tools/dissolve.py:81-223) -------------------------------------------------------------------------


def dissolve_metacells(umis: sp.spmatrix, candidates: np.ndarray, deviants: Optional[np.ndarray], cell_umis: np.ndarray, *,
                       target_metacell_size: int = 48, min_metacell_size: int = 12, target_metacell_umis: int = 160000,
                       min_robust_size_factor: float = 0.5, min_convincing_gene_fold_factor: Optional[float] = 3.0):
    """``(metacell, dissolved)`` per cell, with scipy on the column-major matrix as the reference has it."""
    data = sp.csc_matrix(umis)
    dissolved_of_cells = np.zeros(umis.shape[0], dtype="bool")
    candidate_of_cells = np.copy(candidates)
    if deviants is not None:
        candidate_of_cells[deviants > 0] = -1

    def compress(indices):  # utilities/computation.py:1936-1946
        unique, consecutive = np.unique(indices, return_inverse=True)
        consecutive += min(unique[0], 0)
        return consecutive.astype(indices.dtype)

    candidate_of_cells = compress(candidate_of_cells)
    candidates_count = np.max(candidate_of_cells) + 1
    total_per_gene = np.asarray(data.sum(axis=0)).reshape(-1)
    assert total_per_gene.dtype == np.float32
    fraction_of_genes = total_per_gene / total_per_gene.sum()
    min_robust_size = int(math.floor(target_metacell_size * min_robust_size_factor))
    min_robust_umis = int(math.floor(target_metacell_umis * min_robust_size_factor))
    did_dissolve = False
    for candidate_index in range(candidates_count):
        cell_indices = np.where(candidate_of_cells == candidate_index)[0]
        size = cell_indices.size
        if size < min_metacell_size:
            keep = False
        elif size >= min_robust_size or np.sum(cell_umis[cell_indices]) >= min_robust_umis:
            keep = True
        elif min_convincing_gene_fold_factor is None:
            keep = True
        else:
            data_of_genes = np.asarray(data[cell_indices, :].sum(axis=0)).reshape(-1)
            expected_of_genes = fraction_of_genes * np.sum(data_of_genes)
            expected_of_genes += 1
            data_of_genes += 1
            data_of_genes /= expected_of_genes
            np.log2(data_of_genes, out=data_of_genes)
            keep = bool(np.any(np.abs(data_of_genes) >= min_convincing_gene_fold_factor))
        if not keep:
            dissolved_of_cells[cell_indices] = True
            candidate_of_cells[cell_indices] = -1
            did_dissolve = True
    return (compress(candidate_of_cells) if did_dissolve else candidate_of_cells), dissolved_of_cells


# ---- collect_metacells (pipeline/collect.py:127-232) ----------------------------------------------------------------------


def collect_metacells(umis: sp.csr_matrix, metacell_of_cells: np.ndarray, *, random_seed: int, metacell_geo_mean: bool = True,
                      metacell_umis_regularization: float = 1 / 16.0, zeros_cell_size_quantile: float = 0.1,
                      metacell_groups: bool = False) -> dict:
    """Metacell by metacell as ``_collect_metacell`` does it (dense sub-matrices, numpy / scipy.stats in float64, this
    file's ``downsample_dense`` for the zeros): ``grouped``, ``total_umis``, ``zeros_downsample_umis`` per metacell and
    ``umis_per_gene`` (float32), ``fraction_per_gene`` (float32), ``zeros_per_gene`` (int32) per metacell and gene."""
    import scipy.stats as ss

    metacell_of_cells = np.asarray(metacell_of_cells).copy()
    metacell_of_cells[metacell_of_cells < 0] = -1
    metacells_count = int(np.max(metacell_of_cells)) + 1
    umis_per_cell = np.asarray(umis.sum(axis=1)).reshape(-1).astype("float64")
    results = {"grouped": [], "total_umis": [], "zeros_downsample_umis": [], "umis_per_gene": [], "fraction_per_gene": [],
               "zeros_per_gene": []}
    for metacell_index in range(metacells_count):
        mask = metacell_of_cells == metacell_index
        umis_of_cells = umis_per_cell[mask]
        total_umis = np.sum(umis_of_cells.astype("int64"))
        dense = np.ascontiguousarray(umis[mask, :].toarray(), dtype="float32")
        zeros_downsample_umis = round(np.quantile(umis_of_cells, zeros_cell_size_quantile))
        downsampled = downsample_dense(dense, np.full(dense.shape[0], zeros_downsample_umis, dtype="int32"), random_seed)
        zeros_per_gene = np.sum(downsampled == 0, axis=0).astype("int32")
        umis_per_gene = np.sum(np.asfortranarray(dense), axis=0).astype("float32")
        if not metacell_geo_mean:
            total_downsampled = np.sum(downsampled)
            fraction_per_gene = np.sum(downsampled, axis=0).astype("float32") / total_downsampled
        else:
            fraction_per_cell = dense / umis_of_cells[:, np.newaxis]
            regularization_per_cell = metacell_umis_regularization / umis_of_cells
            weight_per_cell = np.log2(umis_of_cells)
            fraction_per_cell = fraction_per_cell + regularization_per_cell[:, np.newaxis]
            regularization = ss.mstats.gmean(regularization_per_cell, weights=weight_per_cell)
            fraction_per_gene = ss.mstats.gmean(fraction_per_cell.transpose(), axis=1, weights=weight_per_cell)
            fraction_per_gene = np.asarray(fraction_per_gene) - regularization
            fraction_per_gene[umis_per_gene == 0] = 0
            fraction_per_gene[fraction_per_gene < 0] = 0
            fraction_per_gene /= np.sum(fraction_per_gene)
        fraction_per_gene = fraction_per_gene.astype("float32")
        if metacell_groups:
            fraction_per_gene *= total_umis
        for name, value in (("grouped", int(np.sum(mask))), ("total_umis", int(total_umis)),
                            ("zeros_downsample_umis", int(zeros_downsample_umis)), ("umis_per_gene", umis_per_gene),
                            ("fraction_per_gene", fraction_per_gene), ("zeros_per_gene", zeros_per_gene)):
            results[name].append(value)
    return {name: (np.vstack(values) if name.endswith("_per_gene") else np.array(values)) for name, values in results.items()}
