"""mcb_dev_rank_matrix_noalloc_f32 (no counterpart in the reference): the
ranks of mcb_dev_rank_matrix_f32 without any allocation, the scratch in a static device array of the library.  Called
through the C-ABI on device buffers, two calls in a row (the array keeps the last call's contents), compared `==` with the
existing entry point (ranks are integers: the tolerance of the rank tests is exact equality).  This file is collected
last on purpose."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _matrix(rows, columns, seed, hard_rows):
    """Smooth rows (the bucket sort keeps them) and, in `hard_rows`, rows the bucket sort hands to the merge sort through
    the scratch list: constant rows, a handful of distinct values, one wide group of equal values."""
    rng = np.random.default_rng(seed)
    matrix = np.abs(rng.normal(0.2, 0.05, size=(rows, columns))).astype("float32")
    for index, row in enumerate(hard_rows):
        kind = index % 3
        if kind == 0:
            matrix[row] = 0.25
        elif kind == 1:
            matrix[row] = rng.integers(0, 4, size=columns)
        else:
            matrix[row, rng.choice(columns, min(columns, 300), replace=False)] = 0.5
    return matrix


def _rank(library, name, device_matrix, rows, columns, ascending, stream):
    from metacells_b200 import _lib

    function = getattr(library, name)
    function.restype = ctypes.c_int
    _lib.check(function(ctypes.c_void_p(device_matrix.data_ptr()), ctypes.c_int64(rows), ctypes.c_int64(columns),
                        ctypes.c_int64(columns), ctypes.c_int(int(ascending)), ctypes.c_void_p(stream)))


def test_rank_matrix_noalloc_matches_rank_matrix_twice_in_a_row():
    import torch

    from metacells_b200 import _lib

    library = _lib.lib()
    stream = torch.cuda.current_stream().cuda_stream
    # the first call leaves many rows in the scratch list, the second few (and of another shape): nothing may leak over
    cases = [(300, 5000, 1, list(range(0, 300, 2)), False), (120, 16001, 2, [7, 50, 51], True), (64, 900, 3, [], False)]
    for rows, columns, seed, hard_rows, ascending in cases:
        matrix = _matrix(rows, columns, seed, hard_rows)
        expected = torch.from_numpy(matrix).cuda()
        actual = expected.clone()
        _rank(library, "mcb_dev_rank_matrix_f32", expected, rows, columns, ascending, stream)
        _rank(library, "mcb_dev_rank_matrix_noalloc_f32", actual, rows, columns, ascending, stream)
        torch.cuda.synchronize()
        assert torch.equal(actual, expected), (rows, columns)
        # every row is a permutation of 1 .. columns
        assert torch.equal(torch.sort(actual, dim=1).values[0], torch.arange(1, columns + 1, device="cuda", dtype=torch.float32))


def test_rank_matrix_noalloc_rejects_sizes_beyond_its_scratch():
    import torch

    from metacells_b200 import _lib

    library = _lib.lib()
    function = library.mcb_dev_rank_matrix_noalloc_f32
    function.restype = ctypes.c_int
    buffer = torch.zeros(16, device="cuda")
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    too_many_rows = function(ctypes.c_void_p(buffer.data_ptr()), ctypes.c_int64(65537), ctypes.c_int64(4), ctypes.c_int64(4),
                             ctypes.c_int(0), stream)
    assert too_many_rows == 2 and b"at most 65536" in library.mcb_last_error()
    too_wide = function(ctypes.c_void_p(buffer.data_ptr()), ctypes.c_int64(1), ctypes.c_int64(24577), ctypes.c_int64(24577),
                        ctypes.c_int(0), stream)
    assert too_wide == 2 and b"at most 24576" in library.mcb_last_error()
    assert function(ctypes.c_void_p(buffer.data_ptr()), ctypes.c_int64(0), ctypes.c_int64(4), ctypes.c_int64(4), ctypes.c_int(0),
                    stream) == 0
