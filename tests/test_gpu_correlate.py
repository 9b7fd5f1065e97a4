"""GPU parity of correlate_dense / cross_correlate_dense / pairs_correlate_dense.

Tolerance (BASELINE.json north_star): 1e-5 absolute against exact (float64) Pearson and against the
oracle; the reference's own tests use atol=1e-6 against np.corrcoef on their inputs
(tests/test_utilities.py:37-105) and so do we on those inputs.
"""
import numpy as np
import pytest
import scipy.sparse as sp

from oracle import oracle
from synth import synth_umis

pytestmark = pytest.mark.gpu

ATOL = 1e-5


def _xt():
    from metacells_b200 import extensions as xt

    return xt


def _corr(dense):
    out = np.empty((dense.shape[0], dense.shape[0]), dtype="float32")
    name = "correlate_dense_float32_t" if dense.dtype == np.float32 else "correlate_dense_float64_t"
    getattr(_xt(), name)(dense, out)
    return out


def _exact(dense):
    with np.errstate(invalid="ignore", divide="ignore"):
        exact = np.atleast_2d(np.corrcoef(dense.astype("float64")))
    exact[np.isnan(exact)] = 0
    np.fill_diagonal(exact, 1)
    return exact


def _log_fraction_pile(cells, genes, features, seed=123456):
    matrix = synth_umis(cells, genes, seed=seed)
    totals = np.asarray(matrix.sum(axis=1)).reshape(-1)
    samples = np.full(cells, oracle.downsample_samples(totals), dtype="int32")
    matrix.data = oracle.downsample_compressed(matrix.data, matrix.indptr, samples, seed)
    gene_totals = np.asarray(matrix.sum(axis=0)).reshape(-1)
    chosen = np.sort(np.argsort(-gene_totals, kind="stable")[:features])
    return oracle.log_fraction(matrix[:, chosen].toarray())


def test_reference_test_corrcoef_inputs():
    np.random.seed(123456)
    dense = sp.rand(101, 10101, density=0.1, format="csr").toarray()
    exact = np.corrcoef(dense)
    for dtype in ("float64", "float32"):
        result = _corr(np.ascontiguousarray(dense, dtype=dtype))
        assert result.shape == (101, 101)
        assert np.min(np.diag(result)) == 1 and np.max(np.diag(result)) == 1
        assert np.allclose(result, exact, atol=1e-6), (dtype, np.abs(result - exact).max())
    zeros = _corr(np.zeros((101, 64), dtype="float32"))
    assert np.all(np.diag(zeros) == 1)
    np.fill_diagonal(zeros, 0)
    assert zeros.min() == 0 and zeros.max() == 0


def test_reference_test_cross_and_pairs_inputs():
    np.random.seed(123456)
    first = sp.rand(51, 10101, density=0.1, format="csr").toarray()
    second = sp.rand(101, 10101, density=0.1, format="csr").toarray()
    both = np.corrcoef(np.vstack([first, second]))[:51, 51:]
    for dtype in ("float32", "float64"):
        out = np.empty((51, 101), dtype="float32")
        getattr(_xt(), f"cross_correlate_dense_{dtype}_t")(first.astype(dtype), second.astype(dtype), out)
        assert np.allclose(out, both, atol=1e-6), (dtype, np.abs(out - both).max())

    np.random.seed(123456)
    first = sp.rand(100, 100, density=0.1, format="csr").toarray()
    second = sp.rand(100, 100, density=0.1, format="csr").toarray()
    with np.errstate(invalid="ignore", divide="ignore"):
        slow = np.array([np.corrcoef(first[row], second[row])[0, 1] for row in range(100)])
    slow[np.isnan(slow)] = 0
    for dtype in ("float32", "float64"):
        out = np.empty(100, dtype="float32")
        getattr(_xt(), f"pairs_correlate_dense_{dtype}_t")(first.astype(dtype), second.astype(dtype), out)
        assert np.allclose(out, slow, atol=1e-6), dtype


@pytest.mark.parametrize("shape", [(1, 5), (2, 3), (127, 31), (128, 32), (129, 33), (300, 1000), (513, 2049)])
def test_ragged_shapes_match_oracle(shape):
    rng = np.random.default_rng(shape[0] * 1000 + shape[1])
    dense = rng.normal(-2.5, 0.6, size=shape).astype("float32")
    if shape[0] > 2:
        dense[1, :] = 3.25  # a constant row correlates 0 with everything
    result = _corr(dense)
    expected = oracle.correlate_dense(dense)
    assert np.abs(result - expected).max() <= ATOL
    assert np.abs(result - _exact(dense)).max() <= ATOL
    assert np.array_equal(result, result.T)
    assert np.all(np.diag(result) == 1)
    if shape[0] > 2:
        off = np.delete(result[1], 1)
        assert np.all(off == 0)


def test_row_strided_view():
    rng = np.random.default_rng(8)
    wide = rng.normal(size=(200, 700)).astype("float32")
    view = wide[:, :650]
    assert not view.flags.c_contiguous
    assert np.abs(_corr(view) - _exact(view)).max() <= ATOL


@pytest.mark.parametrize("cells,features", [(1500, 1000), (2048, 2000), (1000, 4000)])
def test_log_fraction_pile_matches_oracle_and_exact(cells, features):
    dense = _log_fraction_pile(cells, 6000, features)
    result = _corr(dense)
    exact = _exact(dense)
    error = np.abs(result - exact).max()
    assert error <= ATOL, error
    expected = oracle.correlate_dense(dense)
    assert np.abs(result - expected).max() <= ATOL
    assert np.array_equal(result, result.T)


def test_cross_matches_symmetric_on_same_input():
    dense = _log_fraction_pile(700, 5000, 1500, seed=5)
    out = np.empty((700, 700), dtype="float32")
    _xt().cross_correlate_dense_float32_t(dense, dense, out)
    symmetric = _corr(dense)
    np.fill_diagonal(out, 1)
    assert np.abs(out - symmetric).max() <= 2e-6
    first, second = dense[:333], dense[333:]
    out = np.empty((333, 367), dtype="float32")
    _xt().cross_correlate_dense_float32_t(first, second, out)
    assert np.abs(out - oracle.cross_correlate_dense(np.ascontiguousarray(first), np.ascontiguousarray(second))).max() <= ATOL


def test_full_size_properties_16k():
    """BASELINE config size (16k cells x 2k feature genes): size-independent properties."""
    rng = np.random.default_rng(1)
    cells, features = 16000, 2000
    dense = np.log2(rng.poisson(0.4, size=(cells, features)).astype("float32") + np.float32(1 / 7))
    result = _corr(dense)
    assert np.all(np.diag(result) == 1)
    assert result.min() >= -1 and result.max() <= 1
    rows = rng.choice(cells, size=24, replace=False)
    exact_rows = _exact(dense[np.concatenate([rows, np.arange(64)])])
    picked = result[np.ix_(np.concatenate([rows, np.arange(64)]), np.concatenate([rows, np.arange(64)]))]
    assert np.abs(picked - exact_rows).max() <= ATOL
    sample = rng.choice(cells, size=512, replace=False)
    block = result[np.ix_(sample, sample)]
    assert np.array_equal(block, block.T)
    assert np.array_equal(result[:2000, 9000:11000], result[9000:11000, :2000].T)
