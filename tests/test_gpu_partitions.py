"""GPU parity tests of the partition optimiser (metacells_b200/csrc/partitions.cu) through the C-ABI
(`mcb_host_optimize_partitions`, `mcb_host_score_partitions`, `mcb_host_glibc_log_f64`), with the
reference's symbol names (`xt.optimize_partitions`, `xt.score_partitions`).  Bar: `==` on every element of
partition_of_nodes and on the float64 score, against the oracle restatement and the reference's golden outputs."""
import os

import numpy as np
import pytest
import scipy.sparse as sp

from oracle import oracle
from synth import metacell_bounds, random_partition_graph, seeds_count_for, synth_umis

pytestmark = pytest.mark.gpu
SEED = 123456


@pytest.fixture(scope="module")
def xt():
    import metacells_b200.extensions as extensions

    return extensions


def graph_args(outgoing, incoming):
    return (outgoing.data, outgoing.indices, outgoing.indptr, incoming.data, incoming.indices, incoming.indptr)


def test_device_log_is_glibc_log(xt):
    rng = np.random.default_rng(SEED)
    values = np.concatenate([
        rng.random(1_000_000) * 2, np.exp(rng.normal(0, 20, 1_000_000)), 1 + rng.normal(0, 0.03, 1_000_000),
        1e-6 + rng.random(500_000) * rng.random(500_000),
        np.array([1.0, 0.0, np.inf, 5e-324, 1e-310, 2.0, 1e-6, 0.9375, 1.064697265625, -1.0, np.nan])])
    mine = xt.glibc_log(values)
    theirs = oracle.glibc_log(values)
    same = (mine.view("uint64") == theirs.view("uint64")) | (np.isnan(mine) & np.isnan(theirs))
    assert same.all(), (values[~same][:5], mine[~same][:5], theirs[~same][:5])


def test_golden_pile_graph(xt, golden_dir):
    fixture = np.load(os.path.join(golden_dir, "partitions.npz"))
    nodes = len(fixture["indptr"]) - 1
    outgoing = sp.csr_matrix((fixture["data"], fixture["indices"], fixture["indptr"]), shape=(nodes, nodes))
    incoming = sp.csc_matrix(outgoing)
    incoming.sort_indices()
    bounds = tuple(fixture["bounds"])
    for name in "abc":
        seed, cold, cooldown_pass, cooldown_node = fixture[f"{name}_params"]
        mask = fixture[f"{name}_mask"]
        partition = fixture["seeds"].copy()
        score = xt.optimize_partitions(*graph_args(outgoing, incoming), int(seed), fixture["cell_umis"], *bounds, cooldown_pass,
                                       cooldown_node, partition, mask, cold)
        assert np.array_equal(partition, fixture[f"{name}_partition"]), name
        assert score == fixture[f"{name}_scores"][0], name
        scored = xt.score_partitions(*graph_args(outgoing, incoming), fixture["cell_umis"], *bounds, partition, mask, False)
        assert scored == fixture[f"{name}_scores"][1], name


def test_golden_two_partitions(xt, golden_dir):
    expected = np.load(os.path.join(golden_dir, "partitions_two.npz"))
    outgoing, incoming, masses, initial = random_partition_graph(300, 12, 2, SEED)
    partition = initial.copy()
    score = xt.optimize_partitions(*graph_args(outgoing, incoming), 7, masses, 30000.0, 400000.0, 600000.0, 20.0, 150.0, 250.0,
                                   0.05, 0.25, partition, np.ones(2, dtype="int8"), 0.95)
    assert np.array_equal(partition, expected["partition"])
    assert score == expected["score"][0]


@pytest.mark.parametrize("nodes,degree,partitions,seed", [(200, 8, 5, 1), (400, 16, 9, 2), (257, 6, 3, 3), (1500, 24, 40, 4),
                                                          (33, 3, 2, 5), (3000, 20, 333, 6)])
def test_random_graphs_against_oracle(xt, nodes, degree, partitions, seed):
    outgoing, incoming, masses, initial = random_partition_graph(nodes, degree, partitions, seed)
    bounds = (20000.0, 60000.0, 150000.0, 10.0, 40.0, 90.0)
    mask = np.ones(max(partitions, 2), dtype="int8")
    mask[seed % partitions] = 0
    mine, theirs = initial.copy(), initial.copy()
    stats = np.zeros(4, dtype="int64")
    their_score = oracle.optimize_partitions(outgoing, incoming, seed, masses, *bounds, 0.03, 0.25, theirs, mask, 0.9, stats)
    my_score = xt.optimize_partitions(*graph_args(outgoing, incoming), seed, masses, *bounds, 0.03, 0.25, mine, mask, 0.9)
    assert np.array_equal(xt.last_optimize_stats[:4], stats), (xt.last_optimize_stats, stats)
    assert np.array_equal(mine, theirs)
    assert my_score == their_score
    for with_orphans in (True, False):
        assert xt.score_partitions(*graph_args(outgoing, incoming), masses, *bounds, mine, mask, with_orphans) == \
            oracle.score_partitions(outgoing, incoming, masses, *bounds, theirs, mask, with_orphans)


def test_pile_graph_with_reference_seeds(xt, ref):
    """A 2 000-cell pile through the oracle pipeline, seeds as tools/candidates.py picks them; when oracle/_ref
    travelled to this box the seeds and the expected answer come from the unmodified reference."""
    cells = 2000
    umis = synth_umis(cells, 6000, seed=SEED)
    cell_umis = np.asarray(umis.sum(axis=1)).reshape(-1).astype("float32")
    outgoing = oracle.cpu_pile_knn_graph(umis, 500, 40, SEED, ref=ref)
    outgoing.sort_indices()
    incoming = sp.csc_matrix(outgoing)
    incoming.sort_indices()
    bounds = metacell_bounds(48, 48 * float(np.median(cell_umis)))
    count = seeds_count_for(float(cell_umis.sum()), cells, bounds)
    if ref is not None:
        initial = np.full(cells, -1, dtype="int32")
        count = ref.choose_seeds(*graph_args(outgoing, incoming), SEED, count, 0.8, 1.0, initial)
    else:
        initial = np.random.default_rng(SEED).permutation(np.arange(cells) % count).astype("int32")
    mask = np.ones(max(count, 2), dtype="int8")
    mine, theirs = initial.copy(), initial.copy()
    if ref is not None:
        their_score = ref.optimize_partitions(*graph_args(outgoing, incoming), SEED, cell_umis, *bounds, 0.02, 0.25, theirs, mask, 0.98)
    else:
        their_score = oracle.optimize_partitions(outgoing, incoming, SEED, cell_umis, *bounds, 0.02, 0.25, theirs, mask, 0.98)
    my_score = xt.optimize_partitions(*graph_args(outgoing, incoming), SEED, cell_umis, *bounds, 0.02, 0.25, mine, mask, 0.98)
    assert np.array_equal(mine, theirs)
    assert my_score == their_score


def test_reference_asserts_are_errors(xt):
    outgoing, incoming, masses, initial = random_partition_graph(64, 4, 2, 5)
    mask = np.ones(2, dtype="int8")
    bad = initial.copy()
    bad[3] = -1
    with pytest.raises(xt.McbError, match="partition_index >= 0"):
        xt.optimize_partitions(*graph_args(outgoing, incoming), 1, masses, 1.0, 2.0, 3.0, 1.0, 2.0, 3.0, 0.02, 0.25, bad, mask, 0.9)
    hole = np.where(initial == 1, 2, initial).astype("int32")  # partition 1 is empty
    with pytest.raises(xt.McbError, match="size_of_partitions"):
        xt.optimize_partitions(*graph_args(outgoing, incoming), 1, masses, 1.0, 2.0, 3.0, 1.0, 2.0, 3.0, 0.02, 0.25, hole,
                               np.ones(3, dtype="int8"), 0.9)
    with pytest.raises(xt.McbError, match="low_mass"):
        xt.optimize_partitions(*graph_args(outgoing, incoming), 1, masses, 3.0, 2.0, 1.0, 1.0, 2.0, 3.0, 0.02, 0.25, initial.copy(),
                               mask, 0.9)
    scaled = sp.csr_matrix((outgoing.data * np.float32(0.5), outgoing.indices, outgoing.indptr), shape=outgoing.shape)
    scaled_in = sp.csc_matrix(scaled)
    scaled_in.sort_indices()
    with pytest.raises(xt.McbError, match="sum to 1"):
        xt.optimize_partitions(*graph_args(scaled, scaled_in), 1, masses, 1.0, 2.0, 3.0, 1.0, 2.0, 3.0, 0.02, 0.25, initial.copy(),
                               mask, 0.9)
