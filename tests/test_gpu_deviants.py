"""GPU parity tests of deviant-cell detection (csrc/gaps.cu, metacells_b200/deviants.py) against the REFERENCE's own
`compute_cell_gaps` (oracle/_ref, compiled from gaps.cpp) and the oracle's numpy restatement of tools/deviants.py:241-356.
Bar: `==` on every per-cell gap (float32) of every round and on the final mask, and on every entry of the dense UMIs and
log-fraction matrices."""
import numpy as np
import pytest

from oracle import oracle
from synth import synth_umis

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ref():
    module = oracle.load_ref()
    if module is None or not hasattr(module, "compute_cell_gaps"):
        pytest.skip("oracle/_ref (with gaps.cpp) is not built")
    return module


def typed_candidates(cells, genes, seed, moved_count, rng):
    """Candidates of 48 cells of one cell type of `synth_umis(cells, genes, seed)`, then `moved_count` cells shuffled
    between candidates (the cells a deviants pass should find)."""
    generator = np.random.default_rng(seed)
    types_count = max(4, cells // 400)
    generator.normal(-3.0, 1.6, size=genes)
    generator.normal(0.0, 0.7, size=(types_count, genes))
    types = generator.integers(0, types_count, size=cells)
    order = np.argsort(types, kind="stable")
    candidates = np.empty(cells, dtype="int32")
    candidates[order] = np.arange(cells) // 48
    moved = rng.choice(cells, size=moved_count, replace=False)
    candidates[moved] = candidates[rng.permutation(moved)]
    return candidates


@pytest.mark.parametrize("gap_skip_cells", [1, 2, 3])
def test_compute_cell_gaps_matches_reference(ref, gap_skip_cells):
    """The native call alone, through the drop-in symbol, on matrices full of equal values (few distinct counts, cells of
    equal totals), with deviant and inactive cells, candidates under four cells and an unlisted candidate index."""
    import metacells_b200.extensions as xt

    rng = np.random.default_rng(100 + gap_skip_cells)
    cells, genes, candidates_count = 1500, 300, 40
    umis = rng.poisson(rng.gamma(0.3, 6.0, size=(1, genes)) * rng.choice([0.2, 1.0, 5.0], size=(cells, 1))).astype("float32")
    umis[rng.random(umis.shape) < 0.02] += 40
    totals = np.maximum(umis.sum(axis=1), 1).astype("float32")
    totals = (np.round(totals / 50) * 50 + 50).astype("float32")  # many cells share a total: equal log fractions
    fractions = umis / totals[:, None]
    logs = np.log2(fractions + (1.0 / np.maximum(totals, np.float32(np.quantile(totals, 0.25))))[:, None]).astype("float32")
    candidates = rng.integers(0, candidates_count, size=cells).astype("int32")
    candidates[candidates == 7] = 8  # no cell in candidate 7
    candidates[np.flatnonzero(candidates == 9)[3:]] = 10  # three cells in candidate 9
    deviant = rng.random(cells) < 0.1
    active = rng.random(cells) < 0.05
    active[candidates == 11] = False  # an inactive candidate
    min_gap = rng.choice([0.5, 1.0, 3.0], size=genes).astype("float32")
    mine = rng.random(cells).astype("float32") * np.float32(0.2)  # what is there is kept when it is larger
    theirs = mine.copy()
    arguments = (candidates, deviant, active, min_gap, candidates_count, gap_skip_cells, 3, 0.1, 8.0)
    xt.compute_cell_gaps(umis, fractions, logs, *arguments, mine)
    ref.compute_cell_gaps(umis, fractions, logs, *arguments, theirs)
    assert (theirs > 0.2).sum() > 20
    assert np.array_equal(mine, theirs), np.flatnonzero(mine != theirs)[:10]


def test_gap_inputs_match_numpy():
    """The dense UMIs and log2(fraction + regularization) made on the device == numpy's, including counts beyond the
    per-cell table (patched with numpy's own log2) and a cell without UMIs."""
    from metacells_b200.deviants import GapInputs, regularization_of_cells
    from metacells_b200.pipeline import Pile

    umis = synth_umis(700, 1500, seed=3, depth_scale=4.0).tolil()
    umis[5, 7] = 300
    umis[6, 8] = 70000
    umis[9, :] = 0
    umis = umis.tocsr()
    umis.sort_indices()
    totals = np.asarray(umis.sum(axis=1)).reshape(-1).astype("float32")
    candidates = (np.arange(700) // 50).astype("int32")
    regularization = regularization_of_cells(totals, candidates, 14, 0.25)
    with Pile(umis) as pile, GapInputs(pile, totals, regularization) as inputs:
        assert inputs.exceptions >= 2
        dense, logs = inputs.read()
    expected_dense = umis.toarray().astype("float32")
    with np.errstate(divide="ignore", invalid="ignore"):
        expected_logs = np.log2(expected_dense / totals[:, None] + regularization[:, None])
    assert np.array_equal(dense, expected_dense)
    assert np.array_equal(logs.view("uint32"), expected_logs.view("uint32"))


@pytest.mark.parametrize("cells, genes, depth, moved, arguments", [
    (2000, 4000, 3.0, 60, {}),
    (2000, 4000, 3.0, 200, {"max_cell_fraction": 0.01}),  # too many deviants: the fold bar is raised in eighths
    (3000, 2500, 6.0, 100, {"gap_skip_cells": 0, "min_gene_fold_factor": 2.0, "noisy": True}),
    (1000, 2000, 3.0, 30, {"gap_skip_cells": 2, "max_gap_cells_count": 5, "cells_regularization_quantile": 0.5}),
])
def test_find_deviant_cells_matches_oracle(ref, cells, genes, depth, moved, arguments):
    from metacells_b200.deviants import find_deviant_cells
    from metacells_b200.pipeline import Pile

    arguments = dict(arguments)
    umis = synth_umis(cells, genes, seed=7, depth_scale=depth)
    candidates = typed_candidates(cells, genes, 7, moved, np.random.default_rng(1))
    if arguments.pop("noisy", False):
        arguments["noisy_genes_mask"] = np.random.default_rng(2).random(genes) < 0.1
    mine, theirs = [], []
    with Pile(umis) as pile:
        deviants = find_deviant_cells(pile, candidates, trace=mine, **arguments)
    expected = oracle.gaps_deviant_cells(umis, candidates, ref, trace=theirs, **arguments)
    assert len(mine) == len(theirs)
    for at, (my_gaps, their_gaps) in enumerate(zip(mine, theirs)):
        assert np.array_equal(my_gaps, their_gaps), (at, np.flatnonzero(my_gaps != their_gaps)[:10])
    assert np.array_equal(deviants, expected)
    assert expected.sum() > 0 and len(theirs) >= 2


def test_dissolve_and_the_whole_pile_match_oracle(ref):
    """`dissolve_metacells` on candidates with tiny, small-but-convincing and small-and-bland members == the oracle's scipy
    restatement; then `compute_direct_metacells` (graph -> candidates -> deviants -> dissolve on one uploaded pile) == the
    oracle's steps chained on the same graph."""
    from metacells_b200.deviants import find_deviant_cells
    from metacells_b200.dissolve import dissolve_metacells
    from metacells_b200.pipeline import Pile, compute_direct_metacells, pile_knn_graph_on_device

    cells, genes = 2000, 4000
    umis = synth_umis(cells, genes, seed=7, depth_scale=3.0)
    cell_umis = np.asarray(umis.sum(axis=1)).reshape(-1).astype("float32")
    rng = np.random.default_rng(4)
    candidates = typed_candidates(cells, genes, 7, 60, rng)
    last = int(candidates.max())
    candidates[rng.choice(cells, 5, replace=False)] = last + 1   # under the minimal size
    candidates[np.flatnonzero(candidates == 3)[:20]] = last + 2  # 20 cells of one type: convincing genes
    candidates[rng.choice(np.flatnonzero(candidates <= last), 18, replace=False)] = last + 3  # 18 random cells: bland
    deviants = rng.random(cells) < 0.03
    arguments = dict(target_metacell_umis=int(48 * cell_umis.mean()))
    folds = dict(min_convincing_gene_fold_factor=3.5, **arguments)  # between the two small candidates' largest folds
    with Pile(umis) as pile:
        metacell, dissolved = dissolve_metacells(pile, candidates, deviants, cell_umis, **folds)
        without_folds = dissolve_metacells(pile, candidates, None, cell_umis, min_convincing_gene_fold_factor=None, **arguments)
    expected = oracle.dissolve_metacells(umis, candidates, deviants, cell_umis, **folds)
    assert np.array_equal(metacell, expected[0]) and np.array_equal(dissolved, expected[1])
    assert dissolved[(candidates == last + 1) & ~deviants].all() and dissolved[(candidates == last + 3) & ~deviants].all()
    assert not dissolved[candidates == last + 2].any()
    expected = oracle.dissolve_metacells(umis, candidates, None, cell_umis, min_convincing_gene_fold_factor=None, **arguments)
    assert np.array_equal(without_folds[0], expected[0]) and np.array_equal(without_folds[1], expected[1])

    whole = compute_direct_metacells(umis, features=1000, candidates_arguments=arguments, dissolve_arguments=arguments)
    graph, _ = pile_knn_graph_on_device(umis, features=1000)
    try:
        outgoing, _ = graph.to_scipy()
    finally:
        graph.close()
    candidate = oracle.compute_candidate_metacells(outgoing, cell_umis, random_seed=123456, **arguments)
    deviant = oracle.gaps_deviant_cells(umis, candidate, ref)
    metacell, dissolved = oracle.dissolve_metacells(umis, candidate, deviant, cell_umis, **arguments)
    assert np.array_equal(whole["candidate"], candidate) and np.array_equal(whole["deviant"], deviant)
    assert np.array_equal(whole["metacell"], metacell) and np.array_equal(whole["dissolved"], dissolved)


def test_deviants_and_dissolve_corner_cases(ref):
    """A pile of no more cells than `max_gap_cells_count` has no deviants (tools/deviants.py:260-261); candidates under four
    cells are never looked at (gaps.cpp:62-64); when every candidate is under the minimal size everything is dissolved."""
    from metacells_b200.deviants import find_deviant_cells
    from metacells_b200.dissolve import dissolve_metacells
    from metacells_b200.pipeline import Pile

    tiny = synth_umis(3, 500, seed=5, depth_scale=3.0)
    with Pile(tiny) as pile:
        assert not find_deviant_cells(pile, np.zeros(3, dtype="int32")).any()
    umis = synth_umis(120, 2000, seed=6, depth_scale=3.0)
    cell_umis = np.asarray(umis.sum(axis=1)).reshape(-1).astype("float32")
    candidates = (np.arange(120) // 3).astype("int32")  # forty candidates of three cells
    with Pile(umis) as pile:
        deviants = find_deviant_cells(pile, candidates)
        assert np.array_equal(deviants, oracle.gaps_deviant_cells(umis, candidates, ref)) and not deviants.any()
        metacell, dissolved = dissolve_metacells(pile, candidates, deviants, cell_umis)
    expected = oracle.dissolve_metacells(umis, candidates, deviants, cell_umis)
    assert np.array_equal(metacell, expected[0]) and np.array_equal(dissolved, expected[1])
    assert dissolved.all() and (metacell == -1).all()
