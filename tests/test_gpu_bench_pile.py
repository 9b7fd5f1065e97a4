"""Parity at the BENCHED size, through the BENCHED entry points (task 1): the 16 000-cell x 27 000-gene
pile `bench.py` times (seed 123456, 2 000 feature genes, k from direct.py:251-267 = 333) goes through
mcb_pile_create_host -> _set_max_cell_total -> _downsample -> _top_genes -> _knn_graph (and mcb_graph_seed_partition)
exactly as `bench.py`'s `hot_path` drives them, and every stage is compared with the oracle:

* downsampled values `==` oracle (downsample.cpp:171-192), feature genes `==` numpy, the dense log-fraction operand `==`
  oracle.log_fraction (direct.py:242-249), similarity <= 1e-5 from float64 Pearson and from the unmodified reference
  (`oracle/_ref`, when built), the k-NN graph `==` oracle.knn_graph on the same similarity (tools/knn_graph.py:251-407);
* the digest of that graph is the one `bench.py` checks its timed pile against (tests/golden/bench_pile.json);
* against the `_ref`-driven pipeline (libstdc++'s unstable sort / nth_element: another order inside groups of equal
  values) stage by stage on equal inputs: the reference's ranks are a valid ranking of every row (differences only inside
  groups of exactly equal similarities), both prunes keep the same multiset of values in every band (differences only
  among entries tied at the cut); then the whole chains are compared and the differing edges counted and printed;
* against the whole reference path from the UMIs (its float32 AVX2 correlation): similarities within 1e-5, edge overlap.
"""
import ctypes
import json
import os

import numpy as np
import pytest
import scipy.sparse as sp

from oracle import oracle

pytestmark = pytest.mark.gpu

CELLS, GENES, FEATURES, SEED = 16000, 27000, 2000, 123456
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "bench_pile.json")


def _p(array):
    return ctypes.c_void_p(array.ctypes.data)


@pytest.fixture(scope="module")
def bench_pile():
    """Everything the benched entry points produce for the bench pile, read back stage by stage."""
    import torch

    import bench
    from metacells_b200 import _lib, pipeline, tools as tl

    library = _lib.lib()

    def call(name, *arguments):
        function = getattr(library, name)
        function.restype = ctypes.c_int
        _lib.check(function(*arguments))

    umis, samples, k = bench.make_pile(CELLS, GENES, SEED)
    max_cell_total = bench.make_pile.max_cell_total
    table = pipeline.log_fraction_table(samples)
    pile = ctypes.c_void_p()
    call("mcb_pile_create_host", _p(umis.data), _p(umis.indices), _p(umis.indptr), ctypes.c_int64(CELLS), ctypes.c_int64(GENES),
         ctypes.byref(pile))
    result = {"umis": umis, "samples": samples, "k": k}
    try:
        call("mcb_pile_set_max_cell_total", pile, ctypes.c_uint64(max_cell_total))
        call("mcb_pile_downsample", pile, ctypes.c_int32(samples), ctypes.c_uint64(SEED))
        genes = np.empty(FEATURES, dtype="int32")
        call("mcb_pile_top_genes", pile, ctypes.c_int64(FEATURES), _p(genes))
        down = np.empty(umis.nnz, dtype="float32")
        call("mcb_pile_read_downsampled", pile, _p(down))
        dense = np.empty((CELLS, FEATURES), dtype="float32")
        call("mcb_pile_log_fraction", pile, None, ctypes.c_int64(0), _p(table), ctypes.c_int64(len(table)),
             ctypes.c_float(np.float32(1 / 7)), _p(dense), None)
        similarity = torch.empty((CELLS, CELLS), dtype=torch.float32, device="cuda")
        call("mcb_pile_similarity", pile, None, ctypes.c_int64(0), _p(table), ctypes.c_int64(len(table)),
             ctypes.c_float(np.float32(1 / 7)), ctypes.c_int(1), ctypes.c_void_p(similarity.data_ptr()), None)
        torch.cuda.synchronize()
        graph = ctypes.c_void_p()
        call("mcb_pile_knn_graph", pile, None, ctypes.c_int64(FEATURES), _p(table), ctypes.c_int64(len(table)),
             ctypes.c_float(np.float32(1 / 7)), ctypes.c_int(1), ctypes.c_int64(k), ctypes.c_double(tl.KNN_BALANCED_RANKS_FACTOR),
             ctypes.c_double(tl.KNN_INCOMING_DEGREE_FACTOR), ctypes.c_double(tl.KNN_OUTGOING_DEGREE_FACTOR),
             ctypes.c_int64(tl.KNN_MIN_OUTGOING_DEGREE), None, ctypes.byref(graph))
        handle = tl.KnnGraph(graph)
        weights, pruned = handle.to_scipy()
        result.update(genes=genes, down=down, dense=dense, similarity=similarity.cpu().numpy(), weights=weights, pruned=pruned,
                      used_k=handle.k, graph=handle)
        del similarity
        yield result
        handle.close()
    finally:
        library.mcb_pile_free(pile)


def test_bench_pile_stages_match_oracle(bench_pile):
    umis, samples = bench_pile["umis"], bench_pile["samples"]
    expected_down = oracle.downsample_compressed(umis.data, umis.indptr, np.full(CELLS, samples, dtype="int32"), SEED)
    assert np.array_equal(bench_pile["down"], expected_down), "downsampled pile differs from the oracle"
    down = sp.csr_matrix((expected_down, umis.indices, umis.indptr), shape=umis.shape)
    totals = np.asarray(down.sum(axis=0)).reshape(-1)
    genes = np.sort(np.argsort(-totals, kind="stable")[:FEATURES]).astype("int32")
    assert np.array_equal(bench_pile["genes"], genes), "feature genes differ"
    dense = oracle.log_fraction(down[:, genes].toarray())
    assert np.array_equal(bench_pile["dense"], dense), "log-fraction operand differs from np.log2(x + 1/7)"
    # similarity: 1e-5 absolute (north_star) against float64 Pearson on a sample of rows against all rows
    rows = np.random.default_rng(7).choice(CELLS, 384, replace=False)
    wide = dense.astype("float64")
    wide -= wide.mean(axis=1, keepdims=True)
    wide /= np.sqrt((wide * wide).sum(axis=1, keepdims=True))
    exact = np.abs(wide[rows] @ wide.T)
    exact[np.arange(len(rows)), rows] = 1.0
    error = float(np.abs(bench_pile["similarity"][rows] - exact).max())
    print(f"bench pile: similarity max |error| vs float64 Pearson on {len(rows)} rows: {error:.3e}")
    assert error <= 1e-5


def test_bench_pile_graph_matches_oracle_and_digest(bench_pile):
    """The graph of the fused pile path `==` the oracle's on the same similarity; its digest is the one bench.py checks."""
    import bench

    similarity = bench_pile["similarity"]
    expected_weights, expected_pruned, _, expected_k = oracle.knn_graph(similarity, bench_pile["k"])
    assert bench_pile["used_k"] == expected_k
    for mine, theirs in ((bench_pile["pruned"], expected_pruned), (bench_pile["weights"], expected_weights)):
        assert np.array_equal(mine.indptr, theirs.indptr)
        assert np.array_equal(mine.indices, theirs.indices)
        assert np.array_equal(mine.data, theirs.data)
    weights = bench_pile["weights"]
    digest = bench.graph_digest(weights.data, weights.indices, weights.indptr)
    print(f"bench pile: graph sha256 {digest} ({weights.nnz} edges, k {expected_k})")
    assert digest == bench.graph_digest(expected_weights.data, expected_weights.indices, expected_weights.indptr)
    key = f"{CELLS}x{GENES}x{FEATURES}@{SEED}"
    if os.environ.get("MCB_WRITE_GOLDEN"):
        written = os.path.join(os.path.dirname(GOLDEN), "..", "..", "gpurun_out", "bench_pile.json")
        golden = json.load(open(written)) if os.path.exists(written) else (json.load(open(GOLDEN)) if os.path.exists(GOLDEN) else {})
        golden[key] = digest
        os.makedirs(os.path.dirname(written), exist_ok=True)
        with open(written, "w") as handle:
            json.dump(golden, handle, indent=1)
        return
    with open(GOLDEN) as handle:
        assert json.load(handle).get(key) == digest, "the bench pile's graph digest differs from tests/golden/bench_pile.json"


def test_bench_pile_seeds_and_partition_match_oracle(bench_pile):
    """mcb_graph_seed_partition on the bench pile's graph == oracle.choose_seeds + oracle.optimize_partitions."""
    from metacells_b200 import pipeline

    umis, weights = bench_pile["umis"], bench_pile["weights"]
    cell_umis = np.asarray(umis.sum(axis=1)).reshape(-1).astype("float32")
    bounds = pipeline.metacell_bounds()
    wanted = pipeline.seeds_count_for(float(sum(cell_umis)), CELLS, bounds)
    communities, score, seeds = bench_pile["graph"].seed_partition(cell_umis, bounds, wanted, random_seed=SEED)
    incoming = sp.csc_matrix(weights)
    incoming.sort_indices()
    expected = np.full(CELLS, -1, dtype="int32")
    expected_seeds = oracle.choose_seeds(weights, incoming, SEED, wanted, 0.85, 0.95, expected)
    assert seeds == expected_seeds
    mask = np.ones(max(expected_seeds, 2), dtype="int8")
    expected_score = oracle.optimize_partitions(weights, incoming, SEED, cell_umis, *bounds, 0.02, 0.25, expected, mask, 1 - 0.02)
    assert np.array_equal(communities, expected) and score == expected_score


def _edge_keys(matrix):
    rows = np.repeat(np.arange(matrix.shape[0], dtype=np.int64), np.diff(matrix.indptr))
    return rows * matrix.shape[1] + matrix.indices


def _ref_knn_graph(ref, similarity, k):
    """tools/knn_graph.py:285-407 with every native call going to the unmodified reference (libstdc++'s orders)."""
    size = similarity.shape[0]
    ranks = np.copy(similarity)
    np.fill_diagonal(ranks, np.min(ranks) - 1)
    ref.rank_matrix_float32_t(ranks, False)
    balanced = oracle.balance_ranks(ranks, k, float(np.sqrt(10)))
    incoming_degree, outgoing_degree = min(int(round(k * 3.0)), size - 1), min(int(round(k * 1.0)), size - 1)

    def prune(compressed, top):
        bands = compressed.shape[0]
        out_data = np.empty(bands * top, dtype="float32")
        out_indices = np.empty(bands * top, dtype="int32")
        out_indptr = np.empty(bands + 1, dtype="int32")
        ref.collect_pruned_float32_t_int32_t_int32_t(top, compressed.data, compressed.indices, compressed.indptr, out_data,
                                                     out_indices, out_indptr)
        nnz = out_indptr[-1]
        constructor = sp.csr_matrix if compressed.format == "csr" else sp.csc_matrix
        return constructor((out_data[:nnz], out_indices[:nnz], out_indptr), shape=compressed.shape)

    rows = np.arange(size)
    best = np.asarray(balanced.argmax(axis=1)).reshape(-1)
    preserved = sp.coo_matrix((np.asarray(balanced[rows, best]).reshape(-1), (rows, best)), shape=balanced.shape)
    first = prune(sp.csc_matrix((balanced.data, balanced.indices, balanced.indptr), shape=balanced.shape), incoming_degree)
    transposed = sp.csr_matrix(first)  # the relayout (scipy: same entries, sorted indices)
    transposed.sort_indices()
    second = prune(transposed, outgoing_degree)
    pruned = sp.csr_matrix(second.maximum(preserved).maximum(preserved.transpose()))
    pruned.sort_indices()
    return pruned, ranks, balanced, first, transposed


def _kept_values_by_band(matrix):
    """The values of every band, sorted inside the band (bands keep their order)."""
    bands = np.repeat(np.arange(len(matrix.indptr) - 1), np.diff(matrix.indptr))
    order = np.lexsort((matrix.data, bands))
    return matrix.data[order]


def test_bench_pile_against_reference_orders(bench_pile, ref):
    """The same similarity through the `_ref` natives, stage by stage.  The GPU graph `==` the oracle's (the test above),
    so the oracle's intermediates stand for the GPU's; the reference orders equal values by whatever libstdc++'s unstable
    std::sort / std::nth_element leave (rank.cpp:52-66, prune_per.cpp:33), here they go by position.  So, on equal inputs:
    (1) the reference's ranks are a valid descending ranking of every row (they differ only inside groups of exactly equal
    similarities); (2) (3) each prune keeps, band by band, exactly the same multiset of values (they differ only in which
    of the entries tied at the cut is kept).  Then the two whole chains are compared and the differing edges counted."""
    if ref is None:
        pytest.skip("oracle/_ref is not built")
    ref.set_threads_count(os.cpu_count() or 1)
    similarity, k = bench_pile["similarity"], bench_pile["k"]
    size = similarity.shape[0]
    theirs, ref_ranks, _, _, _ = _ref_knn_graph(ref, similarity, k)
    # (1) ranks
    work = np.copy(similarity)
    np.fill_diagonal(work, np.min(work) - 1)
    oracle_ranks = oracle.rank_matrix(work, ascending=False)
    rows_with_other_ranks = 0
    swapped = 0
    for start in range(0, size, 1024):
        block = work[start:start + 1024]
        descending = -np.sort(-block, axis=1)
        positions = ref_ranks[start:start + 1024].astype(np.int64) - 1
        assert np.array_equal(np.sort(positions, axis=1), np.broadcast_to(np.arange(size), positions.shape)), "not a permutation"
        assert np.array_equal(np.take_along_axis(descending, positions, axis=1), block), "the reference's ranks are not a ranking"
        other = ref_ranks[start:start + 1024] != oracle_ranks[start:start + 1024]
        swapped += int(other.sum())
        rows_with_other_ranks += int(other.any(axis=1).sum())
    print(f"bench pile vs _ref orders: ranks differ at {swapped} of {size * size} entries in {rows_with_other_ranks} rows, all "
          f"inside groups of exactly equal similarities")
    # (2) (3) the prunes on equal inputs
    balanced = oracle.balance_ranks(oracle_ranks, k, float(np.sqrt(10)))
    incoming_degree, outgoing_degree = min(int(round(k * 3.0)), size - 1), min(int(round(k * 1.0)), size - 1)

    def ref_prune(compressed, top):
        bands = compressed.shape[0]
        out_data = np.empty(bands * top, dtype="float32")
        out_indices = np.empty(bands * top, dtype="int32")
        out_indptr = np.empty(bands + 1, dtype="int32")
        ref.collect_pruned_float32_t_int32_t_int32_t(top, compressed.data, compressed.indices, compressed.indptr, out_data,
                                                     out_indices, out_indptr)
        nnz = out_indptr[-1]
        constructor = sp.csr_matrix if compressed.format == "csr" else sp.csc_matrix
        return constructor((out_data[:nnz], out_indices[:nnz], out_indptr), shape=compressed.shape)

    as_columns = sp.csc_matrix((balanced.data, balanced.indices, balanced.indptr), shape=balanced.shape)
    first_mine, first_theirs = oracle.prune_per(as_columns, incoming_degree), ref_prune(as_columns, incoming_degree)
    assert np.array_equal(first_mine.indptr, first_theirs.indptr)
    assert np.array_equal(_kept_values_by_band(first_mine), _kept_values_by_band(first_theirs)), "incoming prune keeps other values"
    other_incoming = int((first_mine.indices != first_theirs.indices).sum())
    transposed = oracle.csc_to_csr(first_mine)
    second_mine, second_theirs = oracle.prune_per(transposed, outgoing_degree), ref_prune(transposed, outgoing_degree)
    assert np.array_equal(second_mine.indptr, second_theirs.indptr)
    assert np.array_equal(_kept_values_by_band(second_mine), _kept_values_by_band(second_theirs)), "outgoing prune keeps other values"
    other_outgoing = int((second_mine.indices != second_theirs.indices).sum())
    print(f"bench pile vs _ref orders: on equal inputs the prunes keep the same values in every band; entries that differ "
          f"(ties at the cut): {other_incoming} of {first_mine.nnz} (incoming), {other_outgoing} of {second_mine.nnz} (outgoing)")
    # the two whole chains
    mine = bench_pile["pruned"]
    mine_keys, their_keys = _edge_keys(mine), _edge_keys(theirs)
    only_mine = np.setdiff1d(mine_keys, their_keys, assume_unique=True)
    only_theirs = np.setdiff1d(their_keys, mine_keys, assume_unique=True)
    common = len(mine_keys) - len(only_mine)
    jaccard = common / (common + len(only_mine) + len(only_theirs))
    print(f"bench pile vs _ref orders, whole chain on the same similarity: {len(mine_keys)} / {len(their_keys)} edges, "
          f"{len(only_mine)} only here, {len(only_theirs)} only in the reference, Jaccard {jaccard:.6f}")
    assert jaccard >= 0.99


def test_bench_pile_against_reference_pipeline(bench_pile, ref):
    """The whole reference path from the UMIs (its own float32 AVX2 correlation): similarities within 1e-5 everywhere;
    the edge sets differ only where ranks of similarities closer than that are swapped -- reported, with the overlap."""
    if ref is None:
        pytest.skip("oracle/_ref is not built")
    ref.set_threads_count(os.cpu_count() or 1)
    dense = bench_pile["dense"]
    their_similarity = np.empty((CELLS, CELLS), dtype="float32")
    ref.correlate_dense_float32_t(dense, their_similarity)
    their_similarity[np.isnan(their_similarity)] = 0.0
    np.fill_diagonal(their_similarity, 1.0)
    np.absolute(their_similarity, out=their_similarity)
    difference = float(np.abs(their_similarity - bench_pile["similarity"]).max())
    print(f"bench pile: similarity max |difference| from the reference's correlate_dense_float32_t: {difference:.3e}")
    assert difference <= 1e-5
    theirs = _ref_knn_graph(ref, their_similarity, bench_pile["k"])[0]
    mine = bench_pile["pruned"]
    mine_keys, their_keys = _edge_keys(mine), _edge_keys(theirs)
    only_mine = np.setdiff1d(mine_keys, their_keys, assume_unique=True)
    only_theirs = np.setdiff1d(their_keys, mine_keys, assume_unique=True)
    common = len(mine_keys) - len(only_mine)
    jaccard = common / (common + len(only_mine) + len(only_theirs))
    # every differing edge: the two similarity matrices order this entry differently against some other entry of one of
    # its two rows, i.e. the row holds another value within the tolerance of it
    size = CELLS
    mine_similarity = bench_pile["similarity"]
    differing = np.concatenate([only_mine, only_theirs])
    far = 0
    ordered_rows = {}
    for key in differing.tolist():
        close = False
        for row, column in ((key // size, key % size), (key % size, key // size)):
            if row not in ordered_rows:
                ordered_rows[row] = np.sort(mine_similarity[row])
            ordered = ordered_rows[row]
            value = mine_similarity[row, column]
            position = np.searchsorted(ordered, value)
            below = ordered[position - 1] if position > 0 else -np.inf
            above = ordered[position + 1] if position + 1 < size else np.inf
            if min(value - below, above - value) <= 2e-5:
                close = True
        far += not close
    print(f"bench pile vs the reference pipeline: {len(only_mine)} edges only here, {len(only_theirs)} only in the reference of "
          f"{len(their_keys)}, Jaccard {jaccard:.6f}; differing edges without a similarity within 2e-5 in either row: {far}")
    assert jaccard >= 0.97


def test_bench_pile_with_selected_genes(ref):
    """bench.py's default path (`--features select`): the same entry points with the feature genes of pipeline/select.py.
    Genes `==` the oracle's selection (scipy sums + the reference's rank_rows on the dense [genes, cells] matrix), graph `==`
    oracle.knn_graph on the similarity of those genes, digest pinned for bench.py."""
    import torch

    import bench
    from metacells_b200 import _lib, pipeline, select, tools as tl

    library = _lib.lib()

    def call(name, *arguments):
        function = getattr(library, name)
        function.restype = ctypes.c_int
        _lib.check(function(*arguments))

    umis, samples, k = bench.make_pile(CELLS, GENES, SEED)
    table = pipeline.log_fraction_table(samples)
    pile = ctypes.c_void_p()
    call("mcb_pile_create_host", _p(umis.data), _p(umis.indices), _p(umis.indptr), ctypes.c_int64(CELLS), ctypes.c_int64(GENES),
         ctypes.byref(pile))
    try:
        call("mcb_pile_set_max_cell_total", pile, ctypes.c_uint64(bench.make_pile.max_cell_total))
        call("mcb_pile_want_gene_statistics", pile, ctypes.c_float(select.SELECT_MIN_GENE_TOP3))
        call("mcb_pile_downsample", pile, ctypes.c_int32(samples), ctypes.c_uint64(SEED))
        view = pipeline.Pile._adopt(pile, (CELLS, GENES))
        view.gene_top_threshold = float(select.SELECT_MIN_GENE_TOP3)
        try:
            genes = np.flatnonzero(select.select_feature_genes(view)).astype("int32")
        finally:
            view._handle = None
        data = np.empty(umis.nnz, dtype="float32")
        call("mcb_pile_read_downsampled", pile, _p(data))
        stored = sp.csr_matrix((data, umis.indices, umis.indptr), shape=umis.shape)
        stored.eliminate_zeros()
        expected_genes = np.flatnonzero(oracle.select_feature_genes(stored, ref=ref)).astype("int32")
        del stored
        print(f"bench pile: {len(genes)} feature genes selected")
        assert np.array_equal(genes, expected_genes)
        similarity = torch.empty((CELLS, CELLS), dtype=torch.float32, device="cuda")
        call("mcb_pile_similarity", pile, _p(genes), ctypes.c_int64(len(genes)), _p(table), ctypes.c_int64(len(table)),
             ctypes.c_float(np.float32(1 / 7)), ctypes.c_int(1), ctypes.c_void_p(similarity.data_ptr()), None)
        torch.cuda.synchronize()
        host_similarity = similarity.cpu().numpy()
        del similarity
        graph = ctypes.c_void_p()
        call("mcb_pile_knn_graph", pile, _p(genes), ctypes.c_int64(len(genes)), _p(table), ctypes.c_int64(len(table)),
             ctypes.c_float(np.float32(1 / 7)), ctypes.c_int(1), ctypes.c_int64(k), ctypes.c_double(tl.KNN_BALANCED_RANKS_FACTOR),
             ctypes.c_double(tl.KNN_INCOMING_DEGREE_FACTOR), ctypes.c_double(tl.KNN_OUTGOING_DEGREE_FACTOR),
             ctypes.c_int64(tl.KNN_MIN_OUTGOING_DEGREE), None, ctypes.byref(graph))
        handle = tl.KnnGraph(graph)
        try:
            weights, pruned = handle.to_scipy()
        finally:
            handle.close()
    finally:
        library.mcb_pile_free(pile)
    expected_weights, expected_pruned, _, _ = oracle.knn_graph(host_similarity, k)
    for mine, theirs in ((pruned, expected_pruned), (weights, expected_weights)):
        assert np.array_equal(mine.indptr, theirs.indptr) and np.array_equal(mine.indices, theirs.indices)
        assert np.array_equal(mine.data, theirs.data)
    digest = bench.graph_digest(weights.data, weights.indices, weights.indptr)
    key = f"{CELLS}x{GENES}xselect@{SEED}"
    print(f"bench pile (selected genes): graph sha256 {digest}")
    if os.environ.get("MCB_WRITE_GOLDEN"):
        written = os.path.join(os.path.dirname(GOLDEN), "..", "..", "gpurun_out", "bench_pile.json")
        golden = json.load(open(written)) if os.path.exists(written) else json.load(open(GOLDEN))
        golden[key] = digest
        os.makedirs(os.path.dirname(written), exist_ok=True)
        with open(written, "w") as handle:
            json.dump(golden, handle, indent=1)
        return
    with open(GOLDEN) as handle:
        assert json.load(handle).get(key) == digest, "the digest of the bench pile's graph (selected genes) is not the pinned one"


def test_bench_pile_to_metacells(ref):
    """The bench pile all the way (`pipeline.compute_direct_metacells`: graph -> the whole candidates loop -> deviant cells
    -> dissolve on one uploaded pile) == the oracle's chain on the same graph: its restatement of tools/candidates.py around
    the pinned natives, the REFERENCE's compute_cell_gaps under tools/deviants.py, scipy's dissolve.  (The oracle's loop
    takes ~45 s of one host core at this size.)"""
    import bench
    from metacells_b200 import pipeline

    if not hasattr(ref, "compute_cell_gaps"):
        pytest.skip("oracle/_ref was built without gaps.cpp")
    umis, _, _ = bench.make_pile(CELLS, GENES, SEED)
    cell_umis = np.asarray(umis.sum(axis=1)).reshape(-1).astype("float32")
    whole = pipeline.compute_direct_metacells(umis, features="select", random_seed=SEED)
    graph, _ = pipeline.pile_knn_graph_on_device(umis, features="select", random_seed=SEED)
    try:
        outgoing, _ = graph.to_scipy()
    finally:
        graph.close()
    with open(GOLDEN) as handle:
        assert bench.graph_digest(outgoing.data, outgoing.indices, outgoing.indptr) == json.load(handle)[
            f"{CELLS}x{GENES}xselect@{SEED}"], "not the graph bench.py times"
    candidate = oracle.compute_candidate_metacells(outgoing, cell_umis, random_seed=SEED)
    assert np.array_equal(whole["candidate"], candidate)
    deviant = oracle.gaps_deviant_cells(umis, candidate, ref)
    assert np.array_equal(whole["deviant"], deviant)
    metacell, dissolved = oracle.dissolve_metacells(umis, candidate, deviant, cell_umis)
    assert np.array_equal(whole["metacell"], metacell) and np.array_equal(whole["dissolved"], dissolved)
    print(f"bench pile: {int(candidate.max()) + 1} candidates, {int(deviant.sum())} deviant cells, "
          f"{int(dissolved.sum())} dissolved cells, {int(metacell.max()) + 1} metacells")
