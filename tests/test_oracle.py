"""CPU tests that pin the oracle (oracle/mc_oracle.c + oracle/oracle.py):
  * against the golden fixtures generated from the compiled, unmodified reference
    (tests/golden/make_golden.py), which travel to machines without /root/reference;
  * against the compiled reference itself (oracle/_ref) on fresh random inputs, when it is present.
"""
import os

import numpy as np
import pytest
import scipy.sparse as sp

from oracle import oracle
from synth import synth_umis


def _golden(golden_dir, name):
    return np.load(os.path.join(golden_dir, name))


def test_downsample_golden(golden_dir):
    import ctypes

    g = _golden(golden_dir, "downsample.npz")
    data, indptr, samples = g["data"], g["indptr"], g["samples"]  # keep the arrays alive across the C call
    for seed in (0, 1, 123456):
        expected = g[f"out_seed_{seed}"]
        out = np.full_like(data, -7.0)
        oracle.lib().mco_downsample_compressed_f32_i32_f32(
            ctypes.c_void_p(data.ctypes.data), ctypes.c_void_p(indptr.ctypes.data), ctypes.c_void_p(out.ctypes.data),
            ctypes.c_void_p(samples.ctypes.data), ctypes.c_int64(len(indptr) - 1), ctypes.c_uint64(seed))
        assert np.array_equal(out, expected), seed


def test_correlate_golden(golden_dir):
    g = _golden(golden_dir, "correlate.npz")
    dense = g["dense"]
    result = oracle.correlate_dense(dense)
    # the oracle restates the reference's float64-accumulating pair kernel: it must sit within 1e-5 of
    # the wheel-equivalent float32 build and much closer to the float64 instantiation
    assert np.abs(result - g["corr"]).max() <= 1e-5
    assert np.abs(result - g["corr64"]).max() <= 2e-7
    assert np.abs(oracle.cross_correlate_dense(np.ascontiguousarray(dense[:40]), np.ascontiguousarray(dense[40:])) - g["cross"]).max() <= 1e-5
    assert np.abs(oracle.pairs_correlate_dense(np.ascontiguousarray(dense[:40]), np.ascontiguousarray(dense[40:80])) - g["pairs"]).max() <= 1e-6


def test_rank_golden(golden_dir):
    g = _golden(golden_dir, "rank.npz")
    distinct = g["distinct"]
    assert np.array_equal(oracle.rank_matrix(distinct, False), g["ranks_desc"])
    assert np.array_equal(oracle.rank_matrix(distinct, True), g["ranks_asc"])
    assert np.array_equal(oracle.rank_rows(distinct, 37), g["kth"])
    indices, data = oracle.collect_top(25, distinct, False)
    assert np.array_equal(indices, g["top_indices"]) and np.array_equal(data, g["top_data"])
    indices, data = oracle.collect_top(25, distinct, True)
    assert np.array_equal(indices, g["top_indices"]) and np.array_equal(data, g["top_rank_data"])


def test_sparse_golden(golden_dir):
    g = _golden(golden_dir, "sparse.npz")
    data, indices, indptr = oracle.collect_pruned(7, g["data"], g["indices"], g["indptr"])
    assert np.array_equal(indptr, g["pruned_indptr"])
    assert np.array_equal(indices, g["pruned_indices"]) and np.array_equal(data, g["pruned_data"])
    data, indices, indptr = oracle.relayout(g["data"], g["indices"], g["indptr"], 60)
    assert np.array_equal(indptr, g["t_indptr"])
    assert np.array_equal(indices, g["t_indices"]) and np.array_equal(data, g["t_data"])


def test_knn_golden(golden_dir):
    g = _golden(golden_dir, "knn.npz")
    assert np.array_equal(oracle.rank_outgoing(g["similarity"]), g["outgoing_ranks"])
    weights, pruned, balanced, k = oracle.knn_graph(g["similarity"], 6)
    assert k == 6
    assert np.allclose(np.asarray(weights.sum(axis=1)).reshape(-1), 1.0, atol=1e-6)
    assert np.diff(pruned.indptr).min() >= 2


def test_reference_property_tests_hold_for_oracle():
    """tests/test_utilities.py:179-222 of the reference, run against the oracle."""
    vector = np.arange(10, dtype="float32")
    out = oracle.downsample_array(vector, 20, 123456)
    assert np.all(out <= vector) and out.sum() == 20
    assert np.array_equal(oracle.downsample_array(vector, 90, 123456), vector)
    dense = np.array([[0, 1, 2], [3, 0, 4], [5, 6, 0], [7, 8, 9]], dtype="float32")
    assert np.allclose(oracle.rank_rows(dense, 1), [1, 3, 5, 8])


# ---- against the compiled reference on fresh inputs -----------------------------------------------


def test_downsample_matches_reference(ref):
    if ref is None:
        pytest.skip("oracle/_ref was not built")
    matrix = synth_umis(64, 2000, seed=3)
    rng = np.random.default_rng(0)
    samples = rng.integers(0, 900, size=64).astype("int32")
    for seed in (0, 7, 123456):
        expected = np.zeros_like(matrix.data)
        ref.downsample_compressed_float32_t_int32_t_float32_t(matrix.data, matrix.indptr, expected, samples, seed)
        assert np.array_equal(oracle.downsample_compressed(matrix.data, matrix.indptr, samples, seed), expected)
    dense = rng.poisson(1.5, size=(9, 257)).astype("float32")
    samples = rng.integers(0, 500, size=9).astype("int32")
    expected = np.zeros_like(dense)
    ref.downsample_dense_float32_t_float32_t(dense, expected, samples, 5)
    assert np.array_equal(oracle.downsample_dense(dense, samples, 5), expected)


def test_selection_matches_reference_modulo_ties(ref):
    """Value-level agreement on inputs WITH ties: the kept values (as multisets) and indptr agree."""
    if ref is None:
        pytest.skip("oracle/_ref was not built")
    rng = np.random.default_rng(1)
    sim = (np.round(rng.random((40, 500)) * 50) / 50).astype("float32")
    ref_indices = np.empty(40 * 30, dtype="int32")
    ref_data = np.empty(40 * 30, dtype="float32")
    ref.collect_top_float32_t(30, sim, ref_indices, ref_data, False)
    indices, data = oracle.collect_top(30, sim, False)
    assert np.array_equal(np.sort(data.reshape(40, 30), axis=1), np.sort(ref_data.reshape(40, 30), axis=1))
    ranks = sim.copy()
    ref.rank_matrix_float32_t(ranks, False)
    mine = oracle.rank_matrix(sim, False)
    # equal up to permutations inside groups of exactly equal values
    for row in range(40):
        order_ref = np.argsort(ranks[row])
        order_mine = np.argsort(mine[row])
        assert np.array_equal(sim[row][order_ref], sim[row][order_mine])


def test_knn_glue_with_reference_natives(ref):
    """The numpy glue of oracle.py with the REFERENCE natives plugged in equals the oracle's own
    natives on a tie-free similarity: pins balance / prune / weigh end to end."""
    if ref is None:
        pytest.skip("oracle/_ref was not built")
    rng = np.random.default_rng(5)
    base = rng.normal(size=(300, 40))
    sim = np.abs(np.corrcoef(base)).astype("float32")
    np.fill_diagonal(sim, 1.0)
    work = sim.copy()
    np.fill_diagonal(work, work.min() - 1)
    ref.rank_matrix_float32_t(work, False)
    assert np.array_equal(work, oracle.rank_outgoing(sim))
    balanced = oracle.balance_ranks(work, 10, float(np.sqrt(10)))
    csc = sp.csc_matrix((balanced.data, balanced.indices, balanced.indptr), shape=balanced.shape)
    out_data = np.empty(300 * 30, dtype="float32")
    out_indices = np.empty(300 * 30, dtype="int32")
    out_indptr = np.empty(301, dtype="int32")
    ref.collect_pruned_float32_t_int32_t_int32_t(30, csc.data, csc.indices, csc.indptr, out_data, out_indices, out_indptr)
    mine = oracle.collect_pruned(30, csc.data, csc.indices, csc.indptr)
    assert np.array_equal(out_indptr, mine[2])
    # balanced ranks are quantised to 1/8, so ties at the cut are expected: compare kept VALUES per band
    for band in range(300):
        lo, hi = out_indptr[band], out_indptr[band + 1]
        assert np.array_equal(np.sort(out_data[lo:hi]), np.sort(mine[0][lo:hi]))
