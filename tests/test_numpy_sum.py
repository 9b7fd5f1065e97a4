"""CPU test of csrc/numpy_sum.cuh through `mcb_host_numpy_add_reduce_f32` (host code, no GPU): the float32 row totals of
`_weigh_edges` (tools/knn_graph.py:399, `ut.sum_per` -> scipy's csr `sum(axis=1)` -> `np.add.reduceat`) in numpy's own
order of additions, `==` numpy and scipy bit for bit -- on values whose float32 sums are not exact, where the order counts
(the graphs of a pile are exact up to k ~ 1000; a grouping pile with k = 1864 is not)."""
import ctypes

import numpy as np
import scipy.sparse as sp

from metacells_b200 import _lib


def library_sums(values, offsets):
    sums = np.empty(len(offsets) - 1, dtype="float32")
    function = _lib.lib().mcb_host_numpy_add_reduce_f32
    function.restype = ctypes.c_int
    _lib.check(function(ctypes.c_void_p(values.ctypes.data), ctypes.c_void_p(offsets.ctypes.data), ctypes.c_int64(len(sums)),
                        ctypes.c_void_p(sums.ctypes.data)))
    return sums


def test_segments_of_every_length_match_numpy():
    rng = np.random.default_rng(0)
    lengths = np.array(list(range(1, 300)) + [511, 512, 513, 1000, 1023, 1024, 1025, 1864, 1865, 2049, 4097, 5000, 9999, 65535])
    offsets = np.concatenate([[0], np.cumsum(lengths)]).astype("int64")
    for make in (lambda n: rng.random(n) * 5000 + 4000, lambda n: rng.normal(size=n), lambda n: np.round(rng.random(n) * 15000) / 8 + 4000):
        values = make(int(offsets[-1])).astype("float32")
        expected = np.add.reduceat(values, offsets[:-1])
        inexact = expected.astype("float64") != np.add.reduceat(values.astype("float64"), offsets[:-1])
        assert inexact.sum() > 5  # the order of the additions is what is being tested (eighths: the long segments only)
        assert np.array_equal(library_sums(values, offsets).view("uint32"), expected.view("uint32"))


def test_row_totals_of_a_wide_graph_match_scipy():
    """Balanced ranks of a graph with k = n - 1 = 1864: multiples of 1/8 between 4 000 and 5 900, 1 864 per row."""
    rng = np.random.default_rng(1)
    size = 1865
    dense = (np.round(rng.random((size, size)) * 15000) / 8 + 4000).astype("float32")
    np.fill_diagonal(dense, 0)
    dense[7, :] = 0  # an empty row
    dense[8, 9:] = 0  # rows of a few entries
    matrix = sp.csr_matrix(dense)
    expected = np.asarray(matrix.sum(axis=1)).reshape(-1).astype("float32")
    mine = library_sums(matrix.data, matrix.indptr.astype("int64"))
    assert np.array_equal(mine.view("uint32"), expected.view("uint32"))
    sequential = np.array([np.cumsum(row.data, dtype="float32")[-1] if row.nnz else 0 for row in matrix])
    assert (sequential != expected).sum() > size // 2  # left to right differs: the case the kernel's second path is for
