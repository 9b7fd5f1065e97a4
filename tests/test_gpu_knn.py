"""GPU parity of rank_matrix / rank_rows / collect_top / collect_pruned / relayout and the fused
k-NN graph builder against the CPU oracle (bit-exact: ranks, indices, indptr and float32 values)."""
import numpy as np
import pytest
import scipy.sparse as sp

from oracle import oracle

pytestmark = pytest.mark.gpu


def _xt():
    from metacells_b200 import extensions as xt

    return xt


def _similarity(cells, seed, ties=True):
    """A symmetric |r|-like matrix in [0, 1] with unit diagonal and (optionally) many exact ties."""
    rng = np.random.default_rng(seed)
    base = rng.random((cells, 24)).astype("float32")
    sim = np.abs(np.corrcoef(base)).astype("float32")
    if ties:
        sim = np.round(sim * 64) / 64
    sim = np.maximum(sim, sim.T).astype("float32")
    np.fill_diagonal(sim, 1.0)
    return sim


@pytest.mark.parametrize("shape,ascending", [((1, 1), True), ((3, 2), False), ((7, 33), True), ((40, 1000), False),
                                             ((5, 4096), True), ((4, 4097), False), ((3, 8192), False), ((3, 8193), True),
                                             ((5, 16384), False), ((2, 16385), True), ((3, 24000), False), ((2, 30000), True)])
def test_rank_matrix_matches_oracle(shape, ascending):
    rng = np.random.default_rng(shape[1])
    matrix = rng.normal(size=shape).astype("float32")
    matrix[:, ::3] = np.round(matrix[:, ::3] * 4) / 4  # exact ties, ordered by position
    if shape[1] > 2:
        matrix[0, 1] = 0.0
        matrix[0, 2] = -0.0
    expected = oracle.rank_matrix(matrix, ascending)
    actual = matrix.copy()
    _xt().rank_matrix_float32_t(actual, ascending)
    assert np.array_equal(actual, expected)


@pytest.mark.parametrize("columns", [3000, 9000, 16001, 20000])
@pytest.mark.parametrize("ascending", [False, True])
def test_rank_matrix_rows_of_every_kind(columns, ascending):
    """The bucket sort takes the rows whose values are spread out; rows with wide groups of ties, constant rows and
    rows with infinite values are passed to the merge sort.  The ranks must not depend on which kernel made them."""
    rng = np.random.default_rng(columns + int(ascending))
    matrix = np.empty((12, columns), dtype="float32")
    matrix[0] = np.abs(rng.normal(0.2, 0.03, size=columns))                    # smooth: the bucket sort
    matrix[0, 5] = matrix[0].min() - 1                                          # the caller's diagonal (knn_graph.py:290)
    matrix[1] = rng.normal(size=columns)
    matrix[2] = 0.25                                                            # constant
    matrix[3] = rng.integers(0, 5, size=columns)                                # five values
    matrix[4] = rng.normal(size=columns)
    matrix[4, rng.choice(columns, 200, replace=False)] = 0.5                    # one wide group inside a smooth row
    matrix[5] = rng.normal(size=columns)
    matrix[5, 7], matrix[5, 11] = np.inf, -np.inf
    matrix[6] = np.where(rng.random(columns) < 0.5, 0.0, rng.normal(size=columns))  # half of the row ties at zero
    matrix[6, 3] = -0.0
    matrix[7] = rng.exponential(size=columns) ** 4                              # a long tail
    matrix[8] = np.arange(columns)[::-1]
    matrix[9] = rng.normal(size=columns) * 1e-30                                # tiny range
    matrix[10] = np.concatenate([rng.normal(-5, 0.001, columns // 2), rng.normal(5, 0.001, columns - columns // 2)])
    matrix[11] = np.round(rng.normal(size=columns) * 64) / 64                   # small groups of ties everywhere
    expected = oracle.rank_matrix(matrix, ascending)
    actual = matrix.copy()
    _xt().rank_matrix_float32_t(actual, ascending)
    for row in range(12):
        assert np.array_equal(actual[row], expected[row]), row


def test_rank_matrix_reference_on_distinct_values(ref):
    """On tie-free rows the result must equal the compiled reference's std::sort ranks."""
    if ref is None:
        pytest.skip("oracle/_ref was not built")
    rng = np.random.default_rng(3)
    matrix = np.stack([rng.permutation(5000) for _ in range(6)]).astype("float32")
    expected = matrix.copy()
    ref.rank_matrix_float32_t(expected, False)
    actual = matrix.copy()
    _xt().rank_matrix_float32_t(actual, False)
    assert np.array_equal(actual, expected)


def test_rank_matrix_row_strided_view():
    rng = np.random.default_rng(4)
    wide = rng.normal(size=(9, 700)).astype("float32")
    view = wide[:, :650]
    expected = oracle.rank_matrix(np.ascontiguousarray(view), False)
    _xt().rank_matrix_float32_t(view, False)
    assert np.array_equal(view, expected)
    assert np.all(np.abs(wide[:, 650:]) < 10)  # untouched


def test_rank_rows_matches_reference_test_and_oracle():
    dense = np.array([[0, 1, 2], [3, 0, 4], [5, 6, 0], [7, 8, 9]], dtype="float32")
    output = np.empty(4, dtype="float32")
    _xt().rank_rows_float32_t(dense, output, 1)
    assert np.allclose(output, [1, 3, 5, 8])  # tests/test_utilities.py:455-465
    rng = np.random.default_rng(6)
    matrix = np.round(rng.normal(size=(50, 3001)) * 8).astype("float32") / 8
    for rank in (0, 1, 1500, 3000):
        output = np.empty(50, dtype="float32")
        _xt().rank_rows_float32_t(matrix, output, rank)
        assert np.array_equal(output, oracle.rank_rows(matrix, rank)), rank
    wide = rng.normal(size=(4, 70000)).astype("float32")
    output = np.empty(4, dtype="float32")
    _xt().rank_rows_float32_t(wide, output, 12345)
    assert np.array_equal(output, np.sort(wide, axis=1)[:, 12345])


@pytest.mark.parametrize("ranks", [False, True])
@pytest.mark.parametrize("cells,degree", [(50, 1), (300, 7), (2000, 150), (2000, 1999)])
def test_collect_top_matches_oracle(cells, degree, ranks):
    sim = _similarity(cells, cells + degree)
    expected_indices, expected_data = oracle.collect_top(degree, sim, ranks)
    indices = np.empty(cells * degree, dtype="int32")
    data = np.empty(cells * degree, dtype="float32")
    _xt().collect_top_float32_t(degree, sim, indices, data, ranks)
    assert np.array_equal(indices, expected_indices)
    assert np.array_equal(data, expected_data)


def _random_csr(rows, cols, density, seed, quantum=8):
    rng = np.random.default_rng(seed)
    matrix = sp.random(rows, cols, density=density, format="csr", random_state=rng, dtype="float32")
    matrix.data = (np.ceil(matrix.data * 100 * quantum) / quantum).astype("float32")  # positive, many ties
    matrix.sort_indices()
    matrix.indices = matrix.indices.astype("int32")
    matrix.indptr = matrix.indptr.astype("int32")
    return matrix


@pytest.mark.parametrize("degree", [0, 1, 5, 40, 10000])
def test_collect_pruned_matches_oracle(degree):
    matrix = _random_csr(300, 2000, 0.02, 9)
    matrix = sp.vstack([matrix, sp.csr_matrix((2, 2000), dtype="float32")], format="csr")  # empty bands
    matrix.indices = matrix.indices.astype("int32")
    matrix.indptr = matrix.indptr.astype("int32")
    expected = oracle.collect_pruned(degree, matrix.data, matrix.indices, matrix.indptr)
    bands = matrix.shape[0]
    out_data = np.empty(max(len(matrix.data), 1), dtype="float32")
    out_indices = np.empty(max(len(matrix.data), 1), dtype="int32")
    out_indptr = np.empty(bands + 1, dtype="int32")
    _xt().collect_pruned_float32_t_int32_t_int32_t(degree, matrix.data, matrix.indices, matrix.indptr, out_data, out_indices, out_indptr)
    nnz = out_indptr[-1]
    assert np.array_equal(out_indptr, expected[2])
    assert np.array_equal(out_indices[:nnz], expected[1])
    assert np.array_equal(out_data[:nnz], expected[0])


def test_collect_pruned_wide_band_not_cached():
    matrix = _random_csr(3, 40000, 0.6, 10)
    expected = oracle.collect_pruned(9000, matrix.data, matrix.indices, matrix.indptr)
    from metacells_b200 import utilities as ut

    pruned = ut.prune_per(matrix, 9000)
    assert np.array_equal(pruned.indptr, expected[2])
    assert np.array_equal(pruned.indices, expected[1])
    assert np.array_equal(pruned.data, expected[0])


def test_relayout_reference_test_and_oracle():
    from metacells_b200 import utilities as ut
    import scipy.stats as stats

    rvs = stats.poisson(10, loc=10).rvs  # tests/test_utilities.py:154-176
    csr = sp.random(20, 20, format="csr", dtype="int32", random_state=123456, data_rvs=rvs).astype("float32")
    csc = ut.relayout_compressed(csr)
    scipy_csc = csr.tocsc()
    assert csc.getformat() == "csc"
    assert np.all(csc.indptr == scipy_csc.indptr)
    assert np.all(csc.toarray() == scipy_csc.toarray())
    back = ut.relayout_compressed(csc)
    assert back.getformat() == "csr"
    assert np.all(back.indptr == csr.indptr) and np.all(back.toarray() == csr.toarray())

    matrix = _random_csr(700, 900, 0.05, 12)
    result = ut.relayout_compressed(matrix)
    data, indices, indptr = oracle.relayout(matrix.data, matrix.indices, matrix.indptr, 900)
    assert np.array_equal(result.indptr, indptr)
    assert np.array_equal(result.indices, indices)
    assert np.array_equal(result.data, data)

    wide = _random_csr(6000, 3, 0.9, 13)  # bands wider than the shared-memory sort capacity
    result = ut.relayout_compressed(wide)
    data, indices, indptr = oracle.relayout(wide.data, wide.indices, wide.indptr, 3)
    assert np.array_equal(result.indices, indices) and np.array_equal(result.data, data)


def _assert_same_csr(actual, expected):
    assert actual.shape == expected.shape
    assert np.array_equal(actual.indptr, expected.indptr)
    assert np.array_equal(actual.indices, expected.indices)
    assert np.array_equal(actual.data, expected.data)


@pytest.mark.parametrize("cells,k,ties", [(64, 5, True), (65, 4, False), (500, 12, True), (1500, 31, True), (1500, 30, False)])
def test_knn_graph_matches_oracle(cells, k, ties):
    from metacells_b200 import tools as tl

    sim = _similarity(cells, cells * 7 + k, ties=ties)
    expected_weights, expected_pruned, _, expected_k = oracle.knn_graph(sim, k)
    weights, pruned, used_k = tl.compute_obs_obs_knn_graph(sim, k=k, return_details=True)
    assert used_k == expected_k
    _assert_same_csr(pruned, expected_pruned)
    _assert_same_csr(weights, expected_weights)
    assert np.allclose(np.asarray(weights.sum(axis=1)).reshape(-1), 1.0, atol=1e-5)


def test_knn_graph_of_every_other_node_matches_oracle():
    """k = n - 1 = 1864 (the grouping pile of a 60 000-cell atlas): rows of 1 864 balanced ranks between 4 000 and 5 900 whose
    float32 totals are not exact, so the weights depend on the order scipy's `sum(axis=1)` adds them in (numpy's pairwise
    order, csrc/numpy_sum.cuh); summed left to right the rows are off by up to 4e-6 and `optimize_partitions` refuses the
    graph ("the outgoing weights of every node sum to 1", partitions.cpp:182-183)."""
    from metacells_b200 import tools as tl

    sim = _similarity(1865, 4, ties=False)
    expected_weights, expected_pruned, _, expected_k = oracle.knn_graph(sim, 1864)
    weights, pruned, used_k = tl.compute_obs_obs_knn_graph(sim, k=1864, return_details=True)
    assert used_k == expected_k == 1864
    _assert_same_csr(pruned, expected_pruned)
    _assert_same_csr(weights, expected_weights)
    totals = np.array([weights.data[weights.indptr[row]:weights.indptr[row + 1]].astype("float64").sum() for row in range(1865)])
    assert np.abs(totals - 1).max() < 1e-6
    rows = np.arange(1865)
    left_to_right = np.array([np.cumsum(expected_pruned.data[expected_pruned.indptr[row]:expected_pruned.indptr[row + 1]],
                                        dtype="float32")[-1] for row in rows])
    assert (left_to_right != np.asarray(expected_pruned.sum(axis=1)).reshape(-1)).sum() > 900  # the order matters here


def test_knn_graph_grows_k_until_min_degree():
    from metacells_b200 import tools as tl

    sim = _similarity(200, 77)
    expected_weights, expected_pruned, _, expected_k = oracle.knn_graph(sim, 1, min_outgoing_degree=6)
    weights, pruned, used_k = tl.compute_obs_obs_knn_graph(sim, k=1, min_outgoing_degree=6, return_details=True)
    assert used_k == expected_k and used_k > 1
    _assert_same_csr(pruned, expected_pruned)
    _assert_same_csr(weights, expected_weights)


def test_knn_graph_full_pile_properties():
    """Pile-size (16k cells) run: size-independent properties of the result."""
    from metacells_b200 import tools as tl

    cells, k = 16000, 333
    rng = np.random.default_rng(2)
    base = rng.normal(size=(cells, 16)).astype("float32")
    sim = np.abs(np.corrcoef(base)).astype("float32")
    np.fill_diagonal(sim, 1.0)
    weights, pruned, used_k = tl.compute_obs_obs_knn_graph(sim, k=k, return_details=True)
    assert used_k == k
    degrees = np.diff(pruned.indptr)
    assert degrees.min() >= 2
    assert np.all(pruned.diagonal() == 0)
    assert pruned.data.min() > 0 and pruned.data.max() <= k * np.sqrt(10) + 1
    assert np.all(np.mod(pruned.data * 8, 1) == 0)  # balanced ranks are quantised to 1/8
    rows = np.repeat(np.arange(cells), degrees)
    assert np.all(np.diff(pruned.indices)[rows[1:] == rows[:-1]] > 0)
    assert np.allclose(np.asarray(weights.sum(axis=1)).reshape(-1), 1.0, atol=1e-5)
    # every row keeps its best balanced edge, in both directions (knn_graph.py:383-388)
    best = np.asarray(pruned.argmax(axis=1)).reshape(-1)
    assert np.all(np.asarray(pruned[best, np.arange(cells)]).reshape(-1) > 0)


def test_pile_feature_genes_selected_on_device():
    """mcb_pile_top_genes == np.sort(np.argsort(-totals, stable)[:count]) (ties by lower gene index), and the
    fused pile path (similarity consumed in place, minimum from the correlate epilogue) == the staged one."""
    from synth import synth_umis

    from metacells_b200 import pipeline, tools as tl

    umis = synth_umis(500, 2500, seed=11)
    with pipeline.Pile(umis) as pile:
        pile.downsample(600, 123456)
        for count in (1, 37, 400, 2499):
            assert np.array_equal(pile.select_top_genes(count), pile.top_total_genes(count)), count
        genes = pile.select_top_genes(400)
        fused = pile.knn_graph(None, k=15)
        weights, _ = fused.to_scipy()
        fused.close()
        staged = pile.knn_graph(genes, k=15)
        staged_weights, _ = staged.to_scipy()
        staged.close()
        down = pile.downsampled()
    assert np.array_equal(weights.indptr, staged_weights.indptr) and np.array_equal(weights.indices, staged_weights.indices)
    assert np.array_equal(weights.data, staged_weights.data)
    dense = oracle.log_fraction(down[:, genes].toarray())
    similarity = tl.compute_obs_obs_similarity(dense)
    separate = tl.compute_obs_obs_knn_graph(similarity, k=15)
    assert np.array_equal(weights.indices, separate.indices) and np.array_equal(weights.data, separate.data)


@pytest.mark.parametrize("pin", [True, False])
def test_piles_one_after_the_other_match_the_single_pile_path(pin):
    """`compute_piles_knn_graphs` starts the upload of the next pile before it asks for the graph of the current one;
    the graphs must be the ones `compute_pile_knn_graph` builds pile by pile (and an empty list yields nothing)."""
    from synth import synth_umis

    from metacells_b200 import pipeline

    piles = [synth_umis(cells, 3000, seed=20 + cells) for cells in (400, 650, 300, 512)]
    assert list(pipeline.compute_piles_knn_graphs([], features=300)) == []
    together = list(pipeline.compute_piles_knn_graphs(iter(piles), features=300, pin=pin))
    assert len(together) == len(piles)
    for umis, (weights, details) in zip(piles, together):
        alone, alone_details = pipeline.compute_pile_knn_graph(umis, features=300)
        assert details["k"] == alone_details["k"] and details["samples"] == alone_details["samples"]
        assert np.array_equal(details["feature_genes"], alone_details["feature_genes"])
        assert np.array_equal(weights.indptr, alone.indptr) and np.array_equal(weights.indices, alone.indices)
        assert np.array_equal(weights.data, alone.data)


def test_piles_walked_by_several_threads_match_the_single_pile_path():
    """`compute_piles_knn_graphs(workers=3)`: three host threads, each pile on its thread's own stream, the same graphs
    in the same order as pile by pile."""
    from synth import synth_umis

    from metacells_b200 import pipeline

    piles = [synth_umis(cells, 3000, seed=40 + cells) for cells in (400, 650, 300, 512, 380, 700, 450)]
    together = list(pipeline.compute_piles_knn_graphs(iter(piles), features=300, workers=3))
    assert len(together) == len(piles)
    for umis, (weights, details) in zip(piles, together):
        alone, alone_details = pipeline.compute_pile_knn_graph(umis, features=300)
        assert details["k"] == alone_details["k"] and details["samples"] == alone_details["samples"]
        assert np.array_equal(details["feature_genes"], alone_details["feature_genes"])
        assert np.array_equal(weights.indptr, alone.indptr) and np.array_equal(weights.indices, alone.indices)
        assert np.array_equal(weights.data, alone.data)


@pytest.mark.parametrize("cells,k,factors,kind", [
    (300, 2, (3.0, 1.0), "ties"), (300, 1, (1.0, 1.0), "ties"), (257, 3, (1.5, 0.5), "ties"), (700, 9, (1.0, 0.34), "smooth"),
    (400, 6, (2.0, 1.0), "hub"), (400, 2, (1.0, 1.0), "hub"), (1200, 40, (1.0, 0.25), "ties"), (2100, 90, (3.0, 1.0), "smooth"),
    (5000, 400, (3.0, 1.0), "smooth"), (5000, 450, (2.0, 0.5), "ties"), (6000, 700, (2.0, 0.5), "ties"),
])
def test_knn_graph_prune_corner_cases(cells, k, factors, kind):
    """The pruning without the transposition (prune_mirror.cuh) against the oracle where its cases differ from the plain
    one: cuts inside groups of equal values, preserved edges that fail the threshold (tiny k), rows that are the preserved
    column of more rows than a warp holds (a hub: the general path takes over), bands wider than 1024 and than 2048
    entries."""
    from metacells_b200 import tools as tl

    sim = _similarity(cells, 1000 + cells + k, ties=kind != "smooth")
    if kind == "hub":  # every cell's most similar cell is cell 0
        sim[0, 1:] = sim[1:, 0] = np.float32(0.999)
    incoming, outgoing = factors
    expected_weights, expected_pruned, _, expected_k = oracle.knn_graph(sim, k, incoming_degree_factor=incoming,
                                                                        outgoing_degree_factor=outgoing, min_outgoing_degree=1)
    weights, pruned, used_k = tl.compute_obs_obs_knn_graph(sim, k=k, incoming_degree_factor=incoming,
                                                           outgoing_degree_factor=outgoing, min_outgoing_degree=1,
                                                           return_details=True)
    assert used_k == expected_k
    _assert_same_csr(pruned, expected_pruned)
    _assert_same_csr(weights, expected_weights)
