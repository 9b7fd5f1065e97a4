"""GPU parity of downsample_* against the CPU oracle (bit-exact), through the C-ABI host entry points."""
import numpy as np
import pytest
import scipy.sparse as sp

from oracle import oracle
from synth import synth_umis

pytestmark = pytest.mark.gpu


def _xt():
    from metacells_b200 import extensions as xt

    return xt


def _run_compressed(matrix, samples, seed, prefill=None):
    out = np.full_like(matrix.data, -7.0) if prefill is None else prefill.copy()
    _xt().downsample_compressed_float32_t_int32_t_float32_t(matrix.data, matrix.indptr, out, samples, seed)
    return out


@pytest.mark.parametrize("seed", [0, 1, 123456])
def test_compressed_matches_oracle(seed):
    matrix = synth_umis(300, 3000, seed=7)
    totals = np.asarray(matrix.sum(axis=1)).reshape(-1)
    samples = np.full(matrix.shape[0], oracle.downsample_samples(totals), dtype="int32")
    expected = oracle.downsample_compressed(matrix.data, matrix.indptr, samples, seed)
    actual = _run_compressed(matrix, samples, seed)
    assert np.array_equal(actual, expected)
    sums = np.add.reduceat(actual, matrix.indptr[:-1])
    assert np.all(sums <= samples[0])


def test_compressed_edge_rows():
    """Rows with nnz in {0,1,2,3,2^k,2^k+-1}; samples in {-1, 0, < total, = total, > total}."""
    rng = np.random.default_rng(5)
    sizes = [0, 1, 2, 3, 4, 5, 7, 8, 9, 15, 16, 17, 31, 32, 33, 255, 256, 257, 1023, 1024, 1025, 0, 1]
    rows = []
    for size in sizes:
        rows.append(rng.integers(1, 9, size=size).astype("float32"))
    data = np.concatenate(rows).astype("float32")
    indptr = np.concatenate([[0], np.cumsum(sizes)]).astype("int32")
    totals = np.array([row.sum() for row in rows])
    for mode in ["minus", "zero", "less", "equal", "more", "mixed"]:
        if mode == "minus":
            samples = np.full(len(sizes), -1)
        elif mode == "zero":
            samples = np.zeros(len(sizes))
        elif mode == "less":
            samples = np.maximum(totals // 2, 0)
        elif mode == "equal":
            samples = totals
        elif mode == "more":
            samples = totals + 3
        else:
            samples = rng.integers(-1, 40, size=len(sizes))
        samples = samples.astype("int32")
        for seed in (0, 1, 123456):
            expected = np.full_like(data, -7.0)
            lib = oracle.lib()
            import ctypes

            lib.mco_downsample_compressed_f32_i32_f32(
                ctypes.c_void_p(data.ctypes.data), ctypes.c_void_p(indptr.ctypes.data), ctypes.c_void_p(expected.ctypes.data),
                ctypes.c_void_p(samples.ctypes.data), ctypes.c_int64(len(sizes)), ctypes.c_uint64(seed))
            out = np.full_like(data, -7.0)
            _xt().downsample_compressed_float32_t_int32_t_float32_t(data, indptr, out, samples, seed)
            assert np.array_equal(out, expected), (mode, seed)


def test_dense_and_array_match_oracle():
    rng = np.random.default_rng(11)
    dense = rng.poisson(2.0, size=(37, 513)).astype("float32")
    samples = rng.integers(0, 1200, size=37).astype("int32")
    expected = oracle.downsample_dense(dense, samples, 123456)
    out = np.zeros_like(dense)
    _xt().downsample_dense_float32_t_float32_t(dense, out, samples, 123456)
    assert np.array_equal(out, expected)

    vector = rng.poisson(3.0, size=5000).astype("float32")
    for count in (0, 10, 4999, 100000):
        expected = oracle.downsample_array(vector, count, 123456)
        out = np.zeros_like(vector)
        _xt().downsample_array_float32_t_float32_t(vector, out, count, 123456)
        assert np.array_equal(out, expected), count


def test_reference_property_tests():
    """tests/test_utilities.py:179-222 of the reference: row sums equal `samples`, never above input."""
    import scipy.stats as stats

    rvs = stats.poisson(10, loc=10).rvs
    matrix = sp.random(1000, 10000, format="csr", dtype="int32", random_state=123456, data_rvs=rvs)
    data = matrix.data.astype("float32")
    indptr = matrix.indptr.astype("int32")
    sums = np.add.reduceat(data, indptr[:-1])
    samples = np.full(1000, int(sums.min()), dtype="int32")
    out = np.zeros_like(data)
    _xt().downsample_compressed_float32_t_int32_t_float32_t(data, indptr, out, samples, 123456)
    assert np.all(out <= data)
    assert np.all(np.add.reduceat(out, indptr[:-1]) == sums.min())


def test_wide_row_goes_through_global_workspace():
    rng = np.random.default_rng(3)
    dense = rng.poisson(0.7, size=(3, 40000)).astype("float32")
    samples = np.array([5000, 100, 27000], dtype="int32")
    expected = oracle.downsample_dense(dense, samples, 99)
    out = np.zeros_like(dense)
    _xt().downsample_dense_float32_t_float32_t(dense, out, samples, 99)
    assert np.array_equal(out, expected)


def test_host_pile_is_downsampled_in_chunks_under_the_upload():
    """A pile of more than 8M non-zeros is uploaded in eight row chunks, each downsampled as it arrives
    (mcb_pile_create_host + mcb_pile_set_max_cell_total + mcb_pile_downsample): same bytes as the oracle."""
    from synth import synth_umis

    from metacells_b200 import pipeline
    from oracle import oracle

    umis = synth_umis(2600, 27000, seed=5)
    assert umis.nnz >= (8 << 20)
    totals = np.asarray(umis.sum(axis=1)).reshape(-1)
    samples = int(np.quantile(totals, 0.3))
    expected = oracle.downsample_compressed(umis.data, umis.indptr, np.full(umis.shape[0], samples, dtype="int32"), 123456)
    for hint in (float(totals.max()), None):
        with pipeline.Pile(umis, max_cell_total=hint) as pile:
            pile.downsample(samples, 123456)
            assert np.array_equal(pile.downsampled().data, expected)


def test_too_small_cell_total_is_reported():
    from synth import synth_umis

    from metacells_b200 import pipeline
    from metacells_b200._lib import McbError

    umis = synth_umis(200, 27000, seed=6)
    totals = np.asarray(umis.sum(axis=1)).reshape(-1)
    assert totals.max() > 4096  # the smallest bitmap holds 2 048 UMIs
    with pipeline.Pile(umis, max_cell_total=1.0) as pile:
        pile.downsample(int(np.quantile(totals, 0.3)), 1)
        with pytest.raises(McbError, match="more than the maximum"):
            pile.downsampled()
