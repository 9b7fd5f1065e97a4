"""Feature-gene selection (pipeline/select.py:145-220 over tools/high.py): the device statistics + the numpy tail of
metacells_b200/select.py against the oracle's restatement on scipy / the reference's rank_rows (`==` on every mask)."""
import numpy as np
import pytest
import scipy.sparse as sp

from oracle import oracle

pytestmark = pytest.mark.gpu


def _downsampled(umis, samples, seed, top=4):
    from metacells_b200 import pipeline

    pile = pipeline.Pile(umis)
    pile.want_gene_statistics(top)
    pile.downsample(samples, seed)
    down = pile.downsampled()
    down.eliminate_zeros()  # computation.py:892-899
    return pile, down


@pytest.mark.parametrize("cells,genes,seed", [(700, 3000, 1), (2500, 6000, 2)])
def test_gene_sums_and_masks_match_the_oracle(cells, genes, seed, ref):
    from synth import synth_umis

    from metacells_b200 import select

    umis = synth_umis(cells, genes, seed=seed)
    pile, down = _downsampled(umis, 800, 123456)
    with pile:
        sums, sums_squared, top_counts = select.gene_sums(pile)
        columns = sp.csc_matrix(down)
        assert np.array_equal(sums, np.asarray(columns.sum(axis=0)).reshape(-1))
        assert np.array_equal(sums_squared, np.asarray(columns.multiply(columns).sum(axis=0)).reshape(-1))
        assert np.array_equal(top_counts >= 4, oracle.high_topN_genes(down, 3, 4, ref=ref))
        relative = select.relative_variance_of_genes(sums, sums_squared, cells)
        assert np.array_equal(relative, oracle.relative_variance_per_column(down), equal_nan=True)
        for arguments in ({}, {"min_genes": 2500}, {"min_genes": genes}, {"min_gene_relative_variance": None, "min_genes": 10},
                          {"min_gene_total": 20, "min_genes": 900}):
            mine = select.select_feature_genes(pile, **arguments)
            assert np.array_equal(mine, oracle.select_feature_genes(down, ref=ref, **arguments)), arguments
        lateral = np.random.default_rng(3).random(genes) < 0.2
        assert np.array_equal(select.select_feature_genes(pile, lateral_genes_mask=lateral),
                              oracle.select_feature_genes(down, lateral_genes_mask=lateral, ref=ref))


def test_sums_beyond_float32_integers_follow_numpy(ref):
    """Genes whose sum of squares passes 2^24: the float32 result depends on the order numpy adds in; the columns are
    fetched and summed by numpy itself."""
    from metacells_b200 import select

    rng = np.random.default_rng(5)
    cells, genes = 3000, 400
    dense = rng.poisson(0.3, size=(cells, genes)).astype("float32")
    dense[:, 7] = rng.integers(150, 260, size=cells)        # squares sum ~ 1.3e8
    dense[::3, 11] = rng.integers(200, 300, size=len(dense[::3, 11]))  # 1000 cells: squares sum ~ 6e7
    dense[:, 13] = 0
    umis = sp.csr_matrix(dense)
    umis.indices, umis.indptr = umis.indices.astype("int32"), umis.indptr.astype("int32")
    pile, down = _downsampled(umis, 100000, 9)  # samples above every total: the counts stay as they are
    with pile:
        totals, squares, _ = pile.gene_statistics()
        assert squares[7] >= 1 << 24 and squares[11] >= 1 << 24 and squares[13] == 0 and totals[13] == 0
        sums, sums_squared, top_counts = select.gene_sums(pile)
        columns = sp.csc_matrix(down)
        assert np.array_equal(sums, np.asarray(columns.sum(axis=0)).reshape(-1))
        assert np.array_equal(sums_squared, np.asarray(columns.multiply(columns).sum(axis=0)).reshape(-1))
        assert np.array_equal(squares.astype("float64"), (dense.astype("float64") ** 2).sum(axis=0))
        assert np.array_equal(top_counts, (dense >= 4).sum(axis=0))
        values, sizes = pile.gene_columns(np.array([11, 13, 7], dtype="int32"))
        assert sizes.tolist() == [int((dense[:, 11] != 0).sum()), 0, cells]
        assert np.array_equal(values[0, :sizes[0]], dense[dense[:, 11] != 0, 11]) and np.array_equal(values[2], dense[:, 7])
        assert np.array_equal(select.select_feature_genes(pile, min_genes=50), oracle.select_feature_genes(down, min_genes=50, ref=ref))


def test_sliding_median_matches_pandas():
    import pandas as pd

    from metacells_b200 import select

    rng = np.random.default_rng(11)
    for count, window in ((101, 101), (500, 101), (27100, 101), (40, 7), (300, 127)):
        values = rng.normal(size=count).astype("float32")
        values[rng.choice(count, count // 10, replace=False)] = np.float32(0.25)  # ties
        if count > 200:
            values[rng.choice(count, 3, replace=False)] = [-np.inf, np.inf, np.nan]  # -inf: log2 of a zero variance
        expected = pd.Series(values).rolling(window).median().values[window - 1:]
        assert np.array_equal(select._device_window_medians(values, window), expected, equal_nan=True), (count, window)
    assert np.isnan(select._device_window_medians(np.array([1, -np.inf, 3], dtype="float32"), 3)).all()


def test_ordered_sliding_median_matches_the_host_path():
    """`mcb_host_sliding_median_ordered_f32` (gathers, end padding and the way back on the device) == the host's gathers
    around pandas' rolling median, as sliding_window_function does them (computation.py:1866-1897), ties in the order
    included."""
    from test_select_host import pandas_window_medians

    from metacells_b200 import select

    rng = np.random.default_rng(12)
    for count, window in ((27000, 101), (150, 101), (5000, 7), (1000, 127)):
        values = rng.normal(size=count).astype("float32")
        values[rng.choice(count, 3, replace=False)] = [-np.inf, np.inf, np.nan]
        order_by = rng.integers(0, count // 4, size=count).astype("float32")  # many equal keys: numpy's order of them counts
        expected = select._sliding_median(values, order_by, window, window_medians=pandas_window_medians)
        mine = select._sliding_median(values, order_by, window)
        assert mine.dtype == np.float64 and np.array_equal(mine, expected, equal_nan=True), (count, window)
        ahead = select._sliding_median(values, order_by, window, order=select._argsort_meanwhile(order_by))
        assert np.array_equal(ahead, expected, equal_nan=True)


def test_small_host_arrays_through_the_page_locked_ring():
    """The order and values of the sliding median reach the device through `upload_small` (runtime.cu: a page-locked ring of
    1 MB per host thread read by a kernel, not the copy engine): enough calls to wrap the ring many times, from two host
    threads at once, with and without the ring (`MCB_NO_SMALL_UPLOADS` is read once, so the second form is a subprocess)."""
    import os
    import subprocess
    import sys
    from concurrent.futures import ThreadPoolExecutor

    from test_select_host import pandas_window_medians

    from metacells_b200 import select

    def walk(seed):
        rng = np.random.default_rng(seed)
        for count in (27000, 40000, 27000, 150, 35000, 27000, 43000, 27000, 200, 27000):  # 12 bytes each: the ring wraps
            values = rng.normal(size=count).astype("float32")
            order_by = rng.integers(0, max(count // 4, 1), size=count).astype("float32")
            expected = select._sliding_median(values, order_by, 101, window_medians=pandas_window_medians)
            assert np.array_equal(select._sliding_median(values, order_by, 101), expected, equal_nan=True), (seed, count)
        return True

    with ThreadPoolExecutor(max_workers=2) as pool:
        assert all(pool.map(walk, (1, 2)))
    if os.environ.get("MCB_NO_SMALL_UPLOADS") is None:
        code = ("import sys; sys.path.insert(0, 'tests'); import numpy as np; from metacells_b200 import select; "
                "from test_select_host import pandas_window_medians as p; r = np.random.default_rng(3); "
                "v = r.normal(size=27000).astype('float32'); o = r.integers(0, 6000, size=27000).astype('float32'); "
                "assert np.array_equal(select._sliding_median(v, o, 101), select._sliding_median(v, o, 101, window_medians=p))")
        root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
        done = subprocess.run([sys.executable, "-c", code], cwd=root, env=dict(os.environ, MCB_NO_SMALL_UPLOADS="1"),
                              capture_output=True, text=True, timeout=600)
        assert done.returncode == 0, done.stderr[-2000:]


def test_pile_graph_on_selected_genes_matches_the_staged_path():
    """`features="select"` through the pile pipeline == selecting with the oracle and building the graph stage by stage."""
    from synth import synth_umis

    from metacells_b200 import pipeline, tools as tl

    umis = synth_umis(900, 4000, seed=21)
    weights, details = pipeline.compute_pile_knn_graph(umis, features="select", k=25, random_seed=123456)
    totals = np.asarray(umis.sum(axis=1)).reshape(-1)
    samples = oracle.downsample_samples(totals)
    assert details["samples"] == samples
    data = oracle.downsample_compressed(umis.data, umis.indptr, np.full(900, samples, dtype="int32"), 123456)
    down = sp.csr_matrix((data, umis.indices, umis.indptr), shape=umis.shape)
    down.eliminate_zeros()
    genes = np.flatnonzero(oracle.select_feature_genes(down)).astype("int32")
    assert np.array_equal(details["feature_genes"], genes) and 100 <= len(genes) < 4000
    similarity = tl.compute_obs_obs_similarity(oracle.log_fraction(down[:, genes].toarray()))
    expected = tl.compute_obs_obs_knn_graph(similarity, k=25)
    assert np.array_equal(weights.indptr, expected.indptr) and np.array_equal(weights.indices, expected.indices)
    assert np.array_equal(weights.data, expected.data)
