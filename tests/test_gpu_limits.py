"""The hard limits DESIGN.md section 6 states, each one pinned: past a limit the library either still returns the
reference's result through its slower path, or refuses with a message that names the limit -- never a wrong answer."""
import numpy as np
import pytest
import scipy.sparse as sp

from oracle import oracle
from synth import metacell_bounds, random_partition_graph

pytestmark = pytest.mark.gpu


def _xt():
    import metacells_b200.extensions as extensions

    return extensions


def _ring(nodes):
    """Every node points at the next one: the cheapest graph of a given size in the form the natives take."""
    indices = ((np.arange(nodes) + 1) % nodes).astype("int32")
    outgoing = sp.csr_matrix((np.ones(nodes, dtype="float32"), indices, np.arange(nodes + 1, dtype="int32")), shape=(nodes, nodes))
    incoming = sp.csc_matrix(outgoing)
    incoming.sort_indices()
    return outgoing, incoming


def _arguments(outgoing, incoming):
    return (outgoing.data, outgoing.indices, outgoing.indptr, incoming.data, incoming.indices.astype("int32"),
            incoming.indptr.astype("int32"))


def test_optimizer_and_seeds_refuse_more_than_65535_nodes():
    from metacells_b200._lib import McbError

    outgoing, incoming = _ring(65536)
    masses = np.ones(65536, dtype="float32")
    partition = (np.arange(65536) % 2).astype("int32")
    with pytest.raises(McbError, match="65535"):
        _xt().optimize_partitions(*_arguments(outgoing, incoming), 1, masses, *metacell_bounds(48, 48), 0.02, 0.25, partition,
                                  np.ones(2, dtype="int8"), 0.98)
    assert np.array_equal(partition, np.arange(65536) % 2)
    seeds = np.full(65536, -1, dtype="int32")
    with pytest.raises(McbError, match="65535"):
        _xt().choose_seeds(*_arguments(outgoing, incoming), 1, 100, 0.85, 0.95, seeds)


def test_seeds_refuse_more_than_16384_seeds():
    from metacells_b200._lib import McbError

    outgoing, incoming = _ring(40000)
    seeds = np.full(40000, -1, dtype="int32")
    with pytest.raises(McbError, match="16384"):
        _xt().choose_seeds(*_arguments(outgoing, incoming), 1, 16385, 0.85, 0.95, seeds)


@pytest.mark.parametrize("partitions", [541, 900, 1200])
def test_optimizer_beyond_the_partitions_of_full_speculation_depth(partitions):
    """More partitions than 16 evaluation warps have shared memory for (P ~ 540): fewer visits are evaluated side by side,
    the partition, the score and the visit / move counts stay the reference's."""
    nodes = 6 * partitions
    outgoing, incoming, masses, initial = random_partition_graph(nodes, 10, partitions, seed=partitions)
    target = float(masses.sum()) / partitions
    bounds = (float(int(target * 0.5)), float(int(target)), float(int(target * 2) - 1), 2.0, 6.0, 11.0)
    hot = np.ones(partitions, dtype="int8")
    mine, theirs = initial.copy(), initial.copy()
    my_score = _xt().optimize_partitions(*_arguments(outgoing, incoming), 7, masses, *bounds, 0.02, 0.25, mine, hot, 0.98)
    stats = np.zeros(4, dtype="int64")
    their_score = oracle.optimize_partitions(outgoing, incoming, 7, masses, *bounds, 0.02, 0.25, theirs, hot, 0.98, stats)
    assert np.array_equal(mine, theirs) and my_score == their_score
    assert np.array_equal(_xt().last_optimize_stats[:4], stats)


@pytest.mark.parametrize("columns", [32768, 32769, 50000])
def test_collect_top_beyond_the_cached_rows(columns):
    """Rows wider than the 32 768 keys the fast path keeps in shared memory take the streaming kernel: same result."""
    rng = np.random.default_rng(columns)
    similarity = rng.random((3, columns)).astype("float32")
    similarity[:, ::7] = np.float32(0.5)  # ties
    for degree, ranks in ((150, False), (40, True)):
        expected_indices, expected_data = oracle.collect_top(degree, similarity, ranks)
        indices = np.empty(3 * degree, dtype="int32")
        data = np.empty(3 * degree, dtype="float32")
        _xt().collect_top_float32_t(degree, similarity, indices, data, ranks)
        assert np.array_equal(indices, expected_indices) and np.array_equal(data, expected_data)


def test_rank_matrix_refuses_rows_wider_than_65535():
    from metacells_b200._lib import McbError

    matrix = np.zeros((1, 65536), dtype="float32")
    with pytest.raises(McbError, match="65535"):
        _xt().rank_matrix_float32_t(matrix, True)


def test_gene_statistics_refuse_more_than_65535_cells():
    from metacells_b200._lib import McbError
    from metacells_b200.pipeline import Pile

    umis = sp.csr_matrix((np.ones(65536, dtype="float32"), np.zeros(65536, dtype="int32"), np.arange(65537, dtype="int32")),
                         shape=(65536, 4))
    with Pile(umis) as pile, pytest.raises(McbError, match="65535"):
        pile.want_gene_statistics(4.0)
