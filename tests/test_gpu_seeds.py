"""GPU parity tests of the seed chooser (metacells_b200/csrc/seeds.cu) through the C-ABI (`mcb_host_choose_seeds`),
with the reference's symbol name (`xt.choose_seeds`).  Bar: `==` on every element of seed_of_nodes and on the
returned seeds count, against the reference's golden outputs and the oracle restatement."""
import os

import numpy as np
import pytest
import scipy.sparse as sp

from oracle import oracle
from seeds_cases import CASES, case_inputs
from synth import random_partition_graph, synth_umis

pytestmark = pytest.mark.gpu
SEED = 123456


@pytest.fixture(scope="module")
def xt():
    import metacells_b200.extensions as extensions

    return extensions


def graph_args(outgoing, incoming):
    return (outgoing.data, outgoing.indices, outgoing.indptr, incoming.data, incoming.indices, incoming.indptr)


def check_against_oracle(xt, outgoing, incoming, seed, max_seeds, low, high, given=None):
    nodes = outgoing.shape[0]
    mine = np.full(nodes, -1, dtype="int32") if given is None else given.copy()
    theirs = mine.copy()
    my_count = xt.choose_seeds(*graph_args(outgoing, incoming), seed, max_seeds, low, high, mine)
    their_count = oracle.choose_seeds(outgoing, incoming, seed, max_seeds, low, high, theirs)
    assert my_count == their_count
    different = np.flatnonzero(mine != theirs)
    assert different.size == 0, (different[:8], mine[different[:8]], theirs[different[:8]])
    return mine, my_count


def test_reference_golden_cases(xt, golden_dir):
    expected = np.load(os.path.join(golden_dir, "seeds.npz"))
    for name in CASES:
        outgoing, incoming, seed, max_seeds, low, high, seed_of_nodes = case_inputs(name, golden_dir)
        count = xt.choose_seeds(*graph_args(outgoing, incoming), seed, max_seeds, low, high, seed_of_nodes)
        assert count == expected[f"{name}_count"][0], name
        assert np.array_equal(seed_of_nodes, expected[f"{name}_seeds"]), name


def test_random_graphs_against_oracle(xt):
    for nodes, degree, clusters, seed, max_seeds, low, high in (
        (300, 10, 4, 11, 9, 0.85, 0.95), (1200, 24, 10, 12, 40, 0.1, 0.3), (700, 5, 3, 13, 300, 0.85, 0.95),
        (2500, 50, 25, 14, 52, 0.85, 0.95), (5000, 12, 30, 15, 100, 0.0, 1.0), (64, 4, 2, 16, 1, 0.85, 0.95),
    ):
        outgoing, incoming, _, _ = random_partition_graph(nodes, degree, clusters, seed)
        check_against_oracle(xt, outgoing, incoming, seed, max_seeds, low, high)
    # seeds given for some of the nodes, as candidates.py:399-407 does for the cells of a split community
    outgoing, incoming, _, _ = random_partition_graph(1500, 20, 8, 21)
    given = np.full(1500, -1, dtype="int32")
    given[np.random.default_rng(21).choice(1500, 400, replace=False)] = np.arange(400) % 5
    check_against_oracle(xt, outgoing, incoming, 21, 30, 0.85, 0.95, given)
    # every seed given already: only do_complete_seeds runs
    check_against_oracle(xt, outgoing, incoming, 22, 5, 0.85, 0.95, given)


def test_pile_graph_against_oracle(xt):
    # a real pile's k-NN graph (built by the CUDA path): connected sizes with wide ties, bands of ~3k edges
    from metacells_b200 import pipeline

    umis = synth_umis(4000, 5000, seed=SEED)
    outgoing, _ = pipeline.compute_pile_knn_graph(umis, feature_genes=np.arange(0, 5000, 5, dtype="int32"), k=83, random_seed=SEED)
    outgoing = sp.csr_matrix(outgoing)
    outgoing.sort_indices()
    incoming = sp.csc_matrix(outgoing)
    incoming.sort_indices()
    seeds, count = check_against_oracle(xt, outgoing, incoming, SEED, 83, 0.85, 0.95)
    assert count == 83 and seeds.min() == 0
    check_against_oracle(xt, outgoing, incoming, 0, 40, 0.5, 1.0)


def test_weights_of_every_magnitude(xt):
    # float32 weights over 60 binades: the float64 total of a node is not exactly summable, so the ordered
    # sum runs; and equal weights, so the cut of a seed's band falls inside a group of ties
    rng = np.random.default_rng(5)
    outgoing, _, _, _ = random_partition_graph(900, 30, 6, 31)
    outgoing.data = np.exp2(rng.integers(-40, 20, size=outgoing.nnz)).astype("float32") * (1 + rng.integers(0, 4, size=outgoing.nnz)).astype("float32")
    incoming = sp.csc_matrix(outgoing)
    incoming.sort_indices()
    check_against_oracle(xt, outgoing, incoming, 31, 20, 0.85, 0.95)
    check_against_oracle(xt, outgoing, incoming, 32, 12, 0.2, 0.9)


def test_wide_graph_selection_in_global_memory(xt):
    # more candidates than the shared-memory selection buffers hold
    outgoing, incoming, _, _ = random_partition_graph(40000, 6, 50, 41)
    check_against_oracle(xt, outgoing, incoming, 41, 120, 0.85, 0.95)


def test_reference_asserts_are_errors(xt):
    outgoing, incoming, _, _ = random_partition_graph(64, 4, 2, 5)
    seed_of_nodes = np.full(64, -1, dtype="int32")
    with pytest.raises(xt.McbError):  # more seeds than candidates (choose_seeds.cpp:318)
        xt.choose_seeds(*graph_args(outgoing, incoming), 1, 65, 0.85, 0.95, seed_of_nodes)
    with pytest.raises(xt.McbError):  # quantiles out of order (choose_seeds.cpp:296-298)
        xt.choose_seeds(*graph_args(outgoing, incoming), 1, 5, 0.9, 0.5, seed_of_nodes)
    seed_of_nodes[:10] = np.arange(10)
    with pytest.raises(xt.McbError):  # more seeds given than wanted (choose_seeds.cpp:366)
        xt.choose_seeds(*graph_args(outgoing, incoming), 1, 5, 0.85, 0.95, seed_of_nodes)
    with pytest.raises(TypeError):
        xt.choose_seeds(*graph_args(outgoing, incoming), 1, 5, 0.85, 0.95, seed_of_nodes.astype("int64"))


def test_tools_mirror(xt):
    # metacells.tools.choose_seeds / optimize_partitions by their reference names and keyword arguments
    from metacells_b200 import tools as tl
    from synth import metacell_bounds

    outgoing, incoming, masses, _ = random_partition_graph(800, 16, 8, 51)
    seeds = tl.choose_seeds(edge_weights=outgoing, max_seeds_count=15, random_seed=51)
    expected = np.full(800, -1, dtype="int32")
    assert oracle.choose_seeds(outgoing, incoming, 51, 15, 0.85, 0.95, expected) == 15
    assert np.array_equal(seeds, expected)
    bounds = metacell_bounds(50, 50 * float(np.median(masses)))
    mine, theirs = seeds.copy(), seeds.copy()
    score = tl.optimize_partitions(edge_weights=outgoing, community_of_nodes=mine, node_umis=masses,
                                   low_partition_umis=bounds[0], target_partition_umis=bounds[1], high_partition_umis=bounds[2],
                                   low_partition_size=bounds[3], target_partition_size=bounds[4], high_partition_size=bounds[5],
                                   random_seed=51)
    expected_score = oracle.optimize_partitions(outgoing, incoming, 51, masses, *bounds, 0.02, 0.25, theirs,
                                                np.ones(15, dtype="int8"), 0.02)
    assert np.array_equal(mine, theirs) and score == expected_score


def test_piles_side_by_side(xt):
    # seeds + first optimisation of several piles at once (one CTA each, one host thread and stream each):
    # the same partitions as one at a time, and as the oracle
    from metacells_b200 import pipeline

    piles = []
    for seed in (61, 62, 63, 64, 65):
        outgoing, _, masses, _ = random_partition_graph(600 + 100 * (seed - 61), 14, 7, seed)
        piles.append((outgoing, masses))
    arguments = dict(random_seed=SEED, target_metacell_size=40)
    together = pipeline.compute_piles_seed_partitions(piles, workers=5, **arguments)
    alone = pipeline.compute_piles_seed_partitions(piles, workers=1, **arguments)
    for (outgoing, masses), (communities, score, seeds_count), (communities_alone, score_alone, _) in zip(piles, together, alone):
        assert np.array_equal(communities, communities_alone) and score == score_alone
        incoming = sp.csc_matrix(outgoing)
        incoming.sort_indices()
        bounds = pipeline.metacell_bounds(target_metacell_size=40)
        wanted = pipeline.seeds_count_for(float(sum(masses)), outgoing.shape[0], bounds)
        expected = np.full(outgoing.shape[0], -1, dtype="int32")
        assert oracle.choose_seeds(outgoing, incoming, SEED, wanted, 0.85, 0.95, expected) == seeds_count
        expected_score = oracle.optimize_partitions(outgoing, incoming, SEED, masses, *bounds, 0.02, 0.25, expected,
                                                    np.ones(max(seeds_count, 2), dtype="int8"), 0.98)
        assert np.array_equal(communities, expected) and score == expected_score


def test_pile_candidates_without_leaving_the_device(xt):
    # k-NN graph -> seeds -> first optimisation with the graph staying on the device (mcb_graph_seed_partition), several
    # piles side by side: the same partitions as the host-array entry points give on the graph read back
    from metacells_b200 import pipeline

    piles = [synth_umis(cells, 3000, seed=70 + cells) for cells in (500, 640, 420)]
    arguments = dict(target_metacell_size=40, target_metacell_umis=60000)
    together = pipeline.compute_piles_candidates(piles, features=300, random_seed=SEED, workers=3, **arguments)
    assert len(together) == len(piles)
    for umis, (communities, score, seeds_count, k) in zip(piles, together):
        outgoing, details = pipeline.compute_pile_knn_graph(umis, features=300, random_seed=SEED)
        assert details["k"] == k
        cell_umis = np.asarray(umis.sum(axis=1)).reshape(-1).astype("float32")
        expected, expected_score, expected_count = pipeline.compute_pile_seed_partition(outgoing, cell_umis, random_seed=SEED,
                                                                                       **arguments)
        assert seeds_count == expected_count and score == expected_score
        assert np.array_equal(communities, expected)


def test_runtime_helpers_for_side_by_side_calls():
    import ctypes

    from metacells_b200 import _lib

    library = _lib.lib()
    device, count = ctypes.c_int(-1), ctypes.c_int(0)
    _lib.check(library.mcb_current_device(ctypes.byref(device)))
    _lib.check(library.mcb_device_sm_count(ctypes.byref(count)))
    assert device.value >= 0 and count.value >= 100  # a B200 has 148 SMs
    _lib.call("mcb_reserve_scratch", ("i64", 64 << 20))
    with pytest.raises(_lib.McbError):
        _lib.call("mcb_reserve_scratch", ("i64", -1))
