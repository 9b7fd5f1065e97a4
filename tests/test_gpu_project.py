"""GPU tests of the projection front-end (`tools.compute_query_atlas_candidates`, C-ABI
`mcb_host_project_top_f32` / `mcb_dev_ordered_top_f32`): the per-row ordered selection is exact (integer / index
work), the correlations agree with the oracle within 1e-5 and the order may differ only among correlations closer
than that."""
import ctypes

import numpy as np
import pytest

from oracle import oracle

pytestmark = pytest.mark.gpu
SEED = 123456


def ordered_top(values: np.ndarray, k: int, stride: int = 0):
    import torch

    from metacells_b200._lib import check, lib

    rows, columns = values.shape
    stride = max(stride, columns)
    padded = np.full((rows, stride), np.float32(7.0), dtype="float32")  # larger than any value: must never be read
    padded[:, :columns] = values
    device = torch.from_numpy(padded).cuda()
    indices = torch.empty((rows, k), dtype=torch.int32, device="cuda")
    kept = torch.empty((rows, k), dtype=torch.float32, device="cuda")
    check(lib().mcb_dev_ordered_top_f32(ctypes.c_void_p(device.data_ptr()), ctypes.c_int64(rows), ctypes.c_int64(columns),
                                        ctypes.c_int64(stride), ctypes.c_int64(k), ctypes.c_void_p(indices.data_ptr()),
                                        ctypes.c_void_p(kept.data_ptr()), ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)))
    torch.cuda.synchronize()
    return indices.cpu().numpy(), kept.cpu().numpy()


def expected_top(values: np.ndarray, k: int):
    order = np.argsort(-values, axis=1, kind="stable")[:, :k]
    return order.astype("int32"), np.take_along_axis(values, order, axis=1)


@pytest.mark.parametrize("rows,columns,k,stride", [(5, 37, 1, 0), (9, 37, 37, 41), (64, 5000, 50, 0), (33, 5003, 50, 5003),
                                                    (7, 70001, 50, 70004), (4, 70000, 1500, 0), (300, 1024, 64, 0)])
def test_ordered_top_is_exact(rows, columns, k, stride):
    rng = np.random.default_rng(SEED + columns + k)
    values = np.tanh(rng.normal(scale=0.4, size=(rows, columns))).astype("float32")
    indices, kept = ordered_top(values, k, stride)
    expected_indices, expected_values = expected_top(values, k)
    assert np.array_equal(indices, expected_indices)
    assert np.array_equal(kept, expected_values)


def test_ordered_top_breaks_ties_by_index():
    rng = np.random.default_rng(SEED)
    values = (rng.integers(-20, 21, size=(40, 9000)) / np.float32(20)).astype("float32")  # 41 distinct values
    values[3, 17] = -0.0
    indices, kept = ordered_top(values, 333)
    expected_indices, expected_values = expected_top(values, 333)
    assert np.array_equal(indices, expected_indices)
    assert np.array_equal(kept, expected_values)


def test_ordered_top_with_more_ties_than_fit_shared_memory():
    """The threshold bin holds tens of thousands of equal values: the exact path (radix select inside the bin, equal
    keys in index order)."""
    rng = np.random.default_rng(SEED + 1)
    values = np.zeros((6, 50000), dtype="float32")
    values[0, rng.choice(50000, 20, replace=False)] = 0.5        # 20 above, the rest tie at 0
    values[1] = 0.25                                              # everything ties
    values[2] = rng.choice(np.array([0.1, 0.1000001, 0.1000002], dtype="float32"), size=50000)  # one bin, three keys
    values[3, ::2] = -1.0
    values[4] = np.float32(3.0)                                   # outside [-1, 1]: clamped into the last bin
    values[4, 100:140] = np.float32(4.0)
    values[5, rng.choice(50000, 45000, replace=False)] = -0.75
    for k in (1, 50, 700):
        indices, kept = ordered_top(values, k)
        expected_indices, expected_values = expected_top(values, k)
        assert np.array_equal(indices, expected_indices), k
        assert np.array_equal(kept, expected_values), k


def log_fraction_like(rows, genes, seed, types=12):
    rng = np.random.default_rng(seed)
    programs = np.random.default_rng(SEED).normal(size=(types, genes)).astype("float32")
    counts = rng.poisson(np.exp(1.0 + 0.6 * programs[rng.integers(0, types, size=rows)])).astype("float32")
    return np.log2(counts / counts.sum(axis=1, keepdims=True) + np.float32(1e-5)).astype("float32")


@pytest.mark.parametrize("queries,atlas_rows,genes,count", [(300, 2500, 400, 50), (1, 77, 33, 50), (130, 129, 1000, 129)])
def test_candidates_match_the_oracle(queries, atlas_rows, genes, count):
    from metacells_b200 import tools as tl

    query = log_fraction_like(queries, genes, SEED + 2)
    atlas = log_fraction_like(atlas_rows, genes, SEED + 3)
    indices, correlations = tl.compute_query_atlas_candidates(query, atlas, candidates_count=count)
    count = min(count, atlas_rows)
    assert indices.shape == (queries, count) and correlations.shape == (queries, count)
    expected_indices, expected_correlations = oracle.query_atlas_candidates(query, atlas, count)
    full = oracle.cross_correlate_dense(query, atlas)
    assert np.abs(correlations - expected_correlations).max() <= 1e-5          # the j-th largest correlation
    assert np.abs(correlations - np.take_along_axis(full, indices.astype("int64"), axis=1)).max() <= 1e-5
    assert np.all(np.diff(correlations, axis=1) <= 0)                           # most correlated first
    for row in range(queries):
        assert len(set(indices[row])) == count
    differs = indices != expected_indices
    # a different atlas row at some position is a tie within the tolerance of the correlation
    assert np.all(np.abs(np.take_along_axis(full, indices.astype("int64"), axis=1) - expected_correlations)[differs] <= 2e-5)
    assert differs.mean() < 0.02


def test_blocks_of_query_rows_reproduce_the_single_block_result(monkeypatch):
    from metacells_b200 import tools as tl

    query = log_fraction_like(700, 200, SEED + 4)
    atlas = log_fraction_like(1500, 200, SEED + 5)
    whole = tl.compute_query_atlas_candidates(query, atlas)
    monkeypatch.setenv("MCB_PROJECT_BLOCK_BYTES", str(256 * 1500 * 4))         # 256 query rows at a time
    blocks = tl.compute_query_atlas_candidates(query, atlas)
    assert np.array_equal(whole[0], blocks[0]) and np.array_equal(whole[1], blocks[1])


def test_bad_arguments_are_reported():
    from metacells_b200 import tools as tl

    with pytest.raises(ValueError):
        tl.compute_query_atlas_candidates(np.zeros((3, 5), dtype="float32"), np.zeros((4, 6), dtype="float32"))
    with pytest.raises(ValueError):
        tl.compute_query_atlas_candidates(np.zeros((3, 5), dtype="float32"), np.zeros((0, 5), dtype="float32"))
