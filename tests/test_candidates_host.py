"""CPU tests of the host side of the candidates loop and of deviant detection: the product's control flow
(metacells_b200/candidates.py, deviants.py) is run with stand-in sessions whose native calls are the oracle's / the
reference's CPU functions, and compared with the oracle's restatements; the oracle's minimal cut is checked against brute
force and its strength against a literal walk of utilities/computation.py:2408-2449."""
import itertools
import math

import numpy as np
import pytest
import scipy.sparse as sp

from oracle import oracle
from synth import synth_umis
from test_gpu_candidates import assert_same_trace, planted_similarity
from test_gpu_deviants import typed_candidates


def symmetric_weights(rng, count, density):
    weights = rng.random((count, count)).astype("float32")
    weights[rng.random((count, count)) > density] = 0
    weights = np.tril(weights, -1)
    return weights + weights.T


def test_stoer_wagner_finds_a_minimal_cut():
    rng = np.random.default_rng(1)
    for count, density in itertools.product((2, 3, 4, 6, 9), (0.3, 0.6, 1.0)):
        weights = symmetric_weights(rng, count, density)
        side = oracle.stoer_wagner_min_cut(weights)
        assert 0 < side.sum() < count
        cut = weights[np.ix_(~side, side)].astype("float64").sum()
        best = min(weights[np.ix_([v for v in range(count) if not mask >> v & 1], [v for v in range(count) if mask >> v & 1])]
                   .astype("float64").sum() for mask in range(1, 2 ** count - 1))
        assert math.isclose(cut, best, rel_tol=1e-12, abs_tol=1e-12), (count, density, cut, best)


def literal_strength(weights, is_second):
    """utilities/computation.py:2389-2449, edge by edge."""
    count = len(weights)
    cut_total = first_total = second_total = 0.0
    for source in range(count):
        for target in range(source):
            weight = weights[source, target]
            if not weight > 0:
                continue
            if is_second[source] != is_second[target]:
                cut_total += weight
            elif is_second[source]:
                second_total += weight
            else:
                first_total += weight
    second_size = int(is_second.sum())
    first_size = count - second_size
    if cut_total == 0:
        return 0.0
    if not (first_total > 0 and second_total > 0):
        return None
    cut_mean = cut_total / (first_size * second_size)
    first_mean = first_total / ((first_size * (first_size - 1)) / 2)
    second_mean = second_total / ((second_size * (second_size - 1)) / 2)
    return cut_mean / math.sqrt(first_mean * second_mean)


def test_cut_strength_follows_the_reference_arithmetic():
    from metacells_b200.candidates import cut_strength

    rng = np.random.default_rng(2)
    seen = set()
    for count, density in itertools.product((4, 9, 30, 60), (0.1, 0.5, 1.0)):
        weights = symmetric_weights(rng, count, density)
        for side in (oracle.stoer_wagner_min_cut(weights), rng.random(count) < 0.5):
            expected = literal_strength(weights, side)
            mine = oracle.min_cut_strength(weights, side)
            assert type(mine) is type(expected) and mine == expected
            seen.add(type(expected))
            if expected is None or 0 < side.sum() < count:
                lower = np.tril(weights, -1)
                sources, targets = np.nonzero(lower > 0)
                totals = []
                for mask in (~side[sources] & ~side[targets], side[sources] & side[targets], side[sources] != side[targets]):
                    totals.append(np.cumsum(lower[sources, targets][mask], dtype="float32")[-1] if mask.any() else np.float32(0))
                product = cut_strength(count - int(side.sum()), int(side.sum()), *totals)
                assert type(product) is type(expected) and product == expected
    assert type(None) in seen and np.float32 in seen


class HostSession:
    """`CandidatesSession` with the oracle's CPU natives behind it."""

    def __init__(self, graph, cell_umis):
        self.outgoing = graph.outgoing
        self.incoming = oracle.csr_to_csc(graph.outgoing)
        self.cell_umis = cell_umis
        self.size = graph.shape[0]

    def __enter__(self):
        return self

    def __exit__(self, *_):
        pass

    def choose_seeds(self, seed_of_nodes, max_seeds_count, *, random_seed, min_seed_size_quantile, max_seed_size_quantile):
        return oracle.choose_seeds(self.outgoing, self.incoming, random_seed, max_seeds_count, min_seed_size_quantile,
                                   max_seed_size_quantile, seed_of_nodes)

    def optimize_partitions(self, community_of_nodes, hot_communities, bounds, *, cold_temperature, random_seed, cooldown_pass,
                            cooldown_node):
        hot_mask = np.zeros(max(int(community_of_nodes.max()) + 1, 2), dtype="int8")
        hot_mask[list(hot_communities)] = 1
        return oracle.optimize_partitions(self.outgoing, self.incoming, random_seed, self.cell_umis, *bounds, cooldown_pass,
                                          cooldown_node, community_of_nodes, hot_mask, cold_temperature)

    def community_weights(self, groups):
        blocks = []
        for group in groups:
            sliced = self.outgoing[group, :][:, group]
            blocks.append((sliced + sliced.T).toarray())
        return blocks

    def min_cuts(self, groups):
        answers = []
        for weights in self.community_weights(groups):
            side = oracle.stoer_wagner_min_cut(weights)
            answers.append((side, oracle.min_cut_strength(weights, side)))
        return answers


class HostGraph:
    def __init__(self, outgoing):
        self.outgoing = outgoing
        self.shape = outgoing.shape


@pytest.mark.parametrize("must_complete_cover", [False, True])
def test_host_loop_matches_oracle_loop(monkeypatch, must_complete_cover):
    import metacells_b200.candidates as candidates

    monkeypatch.setattr(candidates, "CandidatesSession", HostSession)
    similarity = planted_similarity([30, 30] * 6 + [26, 4] * 8, 3)
    outgoing = oracle.knn_graph(similarity, 30)[0]
    cell_umis = np.random.default_rng(5).integers(800, 1200, size=len(similarity)).astype("float32")
    arguments = dict(random_seed=123456, target_metacell_size=60, target_metacell_umis=60000,
                     must_complete_cover=must_complete_cover)
    mine, theirs = [], []
    candidate = candidates.compute_candidate_metacells(HostGraph(outgoing), cell_umis, trace=mine, **arguments)
    expected = oracle.compute_candidate_metacells(outgoing, cell_umis, trace=theirs, **arguments)
    assert_same_trace(mine, theirs)
    assert np.array_equal(candidate, expected)
    assert sum(step[1] for step in theirs if step[0] == "split") > 0


def test_cancel_and_members_helpers():
    from metacells_b200.candidates import _cancel_communities, _members_of

    communities = np.array([3, -1, 0, 2, 3, 0, -1, 5, 2, 2], dtype="int32")
    members = _members_of(communities)
    assert [group.tolist() for group in members] == [[2, 5], [], [3, 8, 9], [0, 4], [], [7]]
    kept = _cancel_communities(communities, {2, 4})
    assert kept == 4 and communities.tolist() == [2, -1, 0, -1, 2, 0, -1, 3, -1, -1]


def test_host_deviants_loop_matches_oracle(monkeypatch):
    """`find_deviant_cells` with the reference's compute_cell_gaps behind a stand-in for the device matrices."""
    import metacells_b200.deviants as deviants

    ref = oracle.load_ref()
    if ref is None or not hasattr(ref, "compute_cell_gaps"):
        pytest.skip("oracle/_ref (with gaps.cpp) is not built")

    class HostInputs:
        def __init__(self, pile, totals, regularization):
            self.umis = np.ascontiguousarray(pile.umis.toarray(), dtype="float32")
            self.fractions = self.umis / totals[:, np.newaxis]
            self.logs = np.log2(self.fractions + regularization[:, np.newaxis])

        def __enter__(self):
            return self

        def __exit__(self, *_):
            pass

        def compute_cell_gaps(self, candidates, deviant, active, *rest):
            ref.compute_cell_gaps(self.umis, self.fractions, self.logs, candidates, deviant, active, *rest)

    class HostPile:
        def __init__(self, umis):
            self.umis = umis
            self.shape = umis.shape

        def cell_totals(self):
            return np.asarray(self.umis.sum(axis=1)).reshape(-1)

    monkeypatch.setattr(deviants, "GapInputs", HostInputs)
    umis = synth_umis(1500, 3000, seed=7, depth_scale=3.0)
    candidates = typed_candidates(1500, 3000, 7, 120, np.random.default_rng(1))
    for arguments in ({}, {"max_cell_fraction": 0.01, "noisy_genes_mask": np.random.default_rng(2).random(3000) < 0.1}):
        mine, theirs = [], []
        found = deviants.find_deviant_cells(HostPile(umis), candidates, trace=mine, **arguments)
        expected = oracle.gaps_deviant_cells(umis, candidates, ref, trace=theirs, **arguments)
        assert len(mine) == len(theirs) and all(np.array_equal(a, b) for a, b in zip(mine, theirs))
        assert np.array_equal(found, expected) and expected.sum() > 0
