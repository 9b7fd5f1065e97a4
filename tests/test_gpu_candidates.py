"""GPU parity tests of the candidate-metacells loop (metacells_b200/candidates.py over `mcb_candidates_*`,
csrc/api_candidates.cu) against the oracle's restatement of tools/candidates.py:121-648.  Bar: `==` on the community
vector after every step of the loop (seeds, every optimisation, every split / cut round, every re-seeding), on the float64
scores, on the sub-graphs, the sides of the minimal cuts and their float32 strengths."""
import numpy as np
import pytest

from oracle import oracle
from synth import synth_umis

pytestmark = pytest.mark.gpu
SEED = 123456


def planted_similarity(halves, seed, bridge=0.3):
    """Cells in tight groups that come in pairs (`halves`: the sizes, two by two): 0.9 inside a group, `bridge` to the
    paired group, 0.05 elsewhere, plus symmetric noise; rows shuffled."""
    rng = np.random.default_rng(seed)
    half_of = np.repeat(np.arange(len(halves)), halves)
    pair_of = half_of // 2
    count = len(half_of)
    similarity = np.where(half_of[:, None] == half_of[None, :], 0.9,
                          np.where(pair_of[:, None] == pair_of[None, :], bridge, 0.05)).astype("float32")
    noise = rng.random((count, count)).astype("float32") * np.float32(0.04)
    similarity = similarity + noise + noise.T
    np.fill_diagonal(similarity, 1.0)
    order = rng.permutation(count)
    return np.ascontiguousarray(similarity[order][:, order], dtype="float32")


def device_graph(similarity, k):
    import metacells_b200.tools as tl

    return tl._knn_graph_handle(similarity, k, tl.KNN_BALANCED_RANKS_FACTOR, tl.KNN_INCOMING_DEGREE_FACTOR,
                                tl.KNN_OUTGOING_DEGREE_FACTOR, tl.KNN_MIN_OUTGOING_DEGREE)


def assert_same_trace(mine, theirs):
    assert [step[0] for step in mine] == [step[0] for step in theirs]
    for at, (my_step, their_step) in enumerate(zip(mine, theirs)):
        assert my_step[1:-1] == their_step[1:-1], (at, my_step[:-1], their_step[:-1])
        assert np.array_equal(my_step[-1], their_step[-1]), (at, my_step[0], int((my_step[-1] != their_step[-1]).sum()))


@pytest.fixture(scope="module")
def planted():
    similarity = planted_similarity([30, 30] * 6 + [26, 4] * 8, 3)
    graph = device_graph(similarity, 30)
    outgoing, _ = graph.to_scipy()
    expected = oracle.knn_graph(similarity, 30)[0]
    assert (outgoing != expected).nnz == 0 and np.array_equal(outgoing.indptr, expected.indptr)
    cell_umis = np.random.default_rng(5).integers(800, 1200, size=len(similarity)).astype("float32")
    yield graph, outgoing, cell_umis
    graph.close()


def test_community_weights_and_min_cuts(planted):
    from metacells_b200.candidates import CandidatesSession

    graph, outgoing, cell_umis = planted
    rng = np.random.default_rng(11)
    community_of = rng.integers(0, 14, size=outgoing.shape[0])
    community_of[rng.random(len(community_of)) < 0.1] = -1  # some nodes in no listed community
    groups = [np.flatnonzero(community_of == community) for community in range(14)]
    groups.append(np.flatnonzero(community_of == -1)[:2])  # a community of two
    groups[3] = groups[3][:1]  # and one of a single node (weights only)
    with CandidatesSession(graph, cell_umis) as session:
        blocks = session.community_weights(groups)
        cuts = session.min_cuts([group for group in groups if len(group) >= 2])
    listed = 0
    for group, block in zip(groups, blocks):
        sliced = outgoing[group, :][:, group]
        expected = (sliced + sliced.T).toarray()
        assert np.array_equal(block, expected)
        if len(group) < 2:
            continue
        is_second, strength = cuts[listed]
        listed += 1
        their_side = oracle.stoer_wagner_min_cut(expected)
        assert np.array_equal(is_second, their_side)
        their_strength = oracle.min_cut_strength(expected, their_side)
        assert type(strength) is type(their_strength) and strength == their_strength
    assert listed == len(cuts)


@pytest.mark.parametrize("must_complete_cover", [False, True])
def test_loop_with_splits_and_cuts_matches_oracle(planted, must_complete_cover):
    from metacells_b200.candidates import compute_candidate_metacells

    graph, outgoing, cell_umis = planted
    arguments = dict(random_seed=SEED, target_metacell_size=60, target_metacell_umis=60000,
                     must_complete_cover=must_complete_cover)
    mine, theirs = [], []
    candidate = compute_candidate_metacells(graph, cell_umis, trace=mine, **arguments)
    expected = oracle.compute_candidate_metacells(outgoing, cell_umis, trace=theirs, **arguments)
    assert_same_trace(mine, theirs)
    assert np.array_equal(candidate, expected)
    splits = sum(step[1] for step in theirs if step[0] == "split")
    cuts = sum(step[2] for step in theirs if step[0] == "split")
    assert splits > 0 and (cuts > 0 or must_complete_cover), "the case no longer exercises the split and cut actions"
    assert any(step[0] == "reseed" for step in theirs)


def test_loop_with_an_external_min_cut(planted):
    """`min_cut=`: the caller's own minimal cut (igraph's, where it is installed) on the sub-graphs the GPU slices."""
    from metacells_b200.candidates import compute_candidate_metacells

    graph, outgoing, cell_umis = planted

    def host_min_cut(weights):
        side = oracle.stoer_wagner_min_cut(weights)
        return side, oracle.min_cut_strength(weights, side)

    arguments = dict(random_seed=SEED, target_metacell_size=60, target_metacell_umis=60000)
    candidate = compute_candidate_metacells(graph, cell_umis, min_cut=host_min_cut, **arguments)
    assert np.array_equal(candidate, oracle.compute_candidate_metacells(outgoing, cell_umis, **arguments))


@pytest.mark.parametrize("cells, genes, target_umis_of_mean", [(2000, 4000, 48), (2000, 4000, 20), (8000, 6000, 48)])
def test_pile_candidates_match_oracle(cells, genes, target_umis_of_mean):
    """The whole pile: UMIs -> k-NN graph on the device -> candidates, against the oracle's loop on the same graph.
    `target_umis_of_mean` 20: the UMIs target is that of 20 average cells, so the UMIs limits bind before the sizes do."""
    from metacells_b200.candidates import compute_candidate_metacells
    from metacells_b200.pipeline import pile_knn_graph_on_device

    umis = synth_umis(cells, genes, seed=7)
    graph, cell_umis = pile_knn_graph_on_device(umis, features=1000)
    try:
        outgoing, _ = graph.to_scipy()
        arguments = dict(random_seed=SEED, target_metacell_umis=int(target_umis_of_mean * cell_umis.mean()))
        mine, theirs = [], []
        candidate = compute_candidate_metacells(graph, cell_umis, trace=mine, **arguments)
        expected = oracle.compute_candidate_metacells(outgoing, cell_umis, trace=theirs, **arguments)
    finally:
        graph.close()
    assert_same_trace(mine, theirs)
    assert np.array_equal(candidate, expected)
    assert sum(step[0] == "optimize" for step in theirs) >= 2  # at least one round of cancelled communities


@pytest.mark.parametrize("cells", [4, 9, 25, 70])
def test_tiny_piles_behave_as_the_oracle(cells):
    """Piles far under a metacell's size (the leftovers of a divide-and-conquer walk): the loop gives the oracle's
    candidates, or both refuse the input (an assertion of the reference's natives surfaces as McbError / ReferenceAssert)."""
    from metacells_b200._lib import McbError
    from metacells_b200.candidates import compute_candidate_metacells
    from metacells_b200.pipeline import pile_knn_graph_on_device

    umis = synth_umis(cells, 1500, seed=31)
    graph, cell_umis = pile_knn_graph_on_device(umis, features=200)
    try:
        outgoing, _ = graph.to_scipy()
        arguments = dict(random_seed=SEED, target_metacell_umis=int(48 * cell_umis.mean()))
        try:
            expected = oracle.compute_candidate_metacells(outgoing, cell_umis, **arguments)
        except (oracle.ReferenceAssert, AssertionError, ValueError, IndexError) as failure:
            expected = type(failure)
        try:
            candidate = compute_candidate_metacells(graph, cell_umis, **arguments)
        except (McbError, AssertionError, ValueError, IndexError) as failure:
            candidate = type(failure)
    finally:
        graph.close()
    if isinstance(expected, type) or isinstance(candidate, type):
        assert isinstance(expected, type) and isinstance(candidate, type), (expected, candidate)
    else:
        assert np.array_equal(candidate, expected)
