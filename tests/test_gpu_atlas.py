"""The atlas resident in HBM and the divide-and-conquer walk over it (pipeline/divide_and_conquer.py:931-1079, 1250-1309):
a pile gathered on the device must be the pile `ut.slice` would cut out and upload."""
import numpy as np
import pytest
import scipy.sparse as sp

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def small_atlas():
    from synth import synth_umis

    return synth_umis(2600, 3000, seed=77)


def _same_csr(mine, theirs):
    assert np.array_equal(mine.indptr, theirs.indptr) and np.array_equal(mine.indices, theirs.indices)
    assert np.array_equal(mine.data, theirs.data)


@pytest.mark.parametrize("index_dtype", ["int32", "int64"])
def test_gathered_pile_is_the_uploaded_pile(small_atlas, index_dtype):
    from metacells_b200 import divide_and_conquer as dc, pipeline

    umis = small_atlas.copy()
    umis.indices = umis.indices.astype(index_dtype)
    umis.indptr = umis.indptr.astype(index_dtype)
    rng = np.random.default_rng(5)
    with dc.Atlas.from_host(umis) as atlas:
        totals = np.asarray(small_atlas.sum(axis=1)).reshape(-1).astype("float32")
        assert np.array_equal(atlas.cell_umis, totals)
        for count in (1, 2, 700, 2600):
            rows = np.sort(rng.choice(2600, count, replace=False))
            sliced = small_atlas[rows]
            with atlas.pile(rows) as gathered, pipeline.Pile(sliced) as uploaded:
                gathered.downsample(500, 123456)
                uploaded.downsample(500, 123456)
                assert np.array_equal(gathered.gene_totals(), uploaded.gene_totals())
                if count < 3:
                    continue
                assert np.array_equal(gathered.select_top_genes(300), uploaded.select_top_genes(300))
                assert np.array_equal(gathered.log_fraction(None), uploaded.log_fraction(None))
                mine, theirs = gathered.knn_graph(None, k=12), uploaded.knn_graph(None, k=12)
                try:
                    for one, other in zip(mine.to_scipy(), theirs.to_scipy()):
                        _same_csr(one, other)
                finally:
                    mine.close()
                    theirs.close()


def test_rows_in_any_order_and_bad_rows(small_atlas):
    from metacells_b200 import divide_and_conquer as dc, pipeline
    from metacells_b200._lib import McbError

    with dc.Atlas.from_host(small_atlas) as atlas:
        rows = np.random.default_rng(6).permutation(2600)[:400]  # not sorted: the pile keeps the caller's order
        with atlas.pile(rows) as gathered, pipeline.Pile(small_atlas[rows]) as uploaded:
            gathered.downsample(400, 7)
            uploaded.downsample(400, 7)
            genes = np.arange(0, 3000, 7, dtype="int32")
            assert np.array_equal(gathered.log_fraction(genes), uploaded.log_fraction(genes))
        with pytest.raises(McbError, match="out of range"):
            atlas.pile(np.array([0, 2600]))


def test_group_sums_and_dense_pile(small_atlas):
    import torch

    from metacells_b200 import divide_and_conquer as dc, pipeline

    rng = np.random.default_rng(8)
    groups = 40
    group_of_cells = rng.integers(-1, groups, size=2600).astype("int32")
    with dc.Atlas.from_host(small_atlas) as atlas:
        sums = atlas.sum_groups(group_of_cells, groups)
        expected = np.zeros((groups, 3000), dtype="float32")
        dense = small_atlas.toarray()
        for group in range(groups):
            expected[group] = dense[group_of_cells == group].sum(axis=0)
        assert np.array_equal(sums.cpu().numpy(), expected)
        with pipeline.Pile.from_dense_device(sums.data_ptr(), groups, 3000) as mine, pipeline.Pile(sp.csr_matrix(expected)) as theirs:
            mine.set_max_cell_total(float(expected.sum(axis=1).max()))
            mine.downsample(3000, 99)
            theirs.downsample(3000, 99)
            assert np.array_equal(mine.gene_totals(), theirs.gene_totals())
            genes = np.arange(0, 3000, 3, dtype="int32")
            assert np.array_equal(mine.log_fraction(genes), theirs.log_fraction(genes))
        torch.cuda.synchronize()


def test_phase_over_the_atlas_matches_pile_by_pile(small_atlas):
    """`compute_piles_phase` == the graph + seeds + first optimisation of every pile cut out on the host."""
    from metacells_b200 import divide_and_conquer as dc, piles as pl, pipeline

    pile_of_cells = pl.random_piles(2600, 900, random_seed=123456).astype("int32")
    pile_of_cells[:5] = -1  # cells in no pile
    with dc.Atlas.from_host(small_atlas) as atlas:
        times = {}
        phase = dc.compute_piles_phase(atlas, pile_of_cells, features=300, random_seed=123456, times=times)
    assert times["graphs_s"] > 0 and len(phase) == int(pile_of_cells.max()) + 1
    for pile, result in enumerate(phase):
        rows = np.flatnonzero(pile_of_cells == pile)
        assert np.array_equal(result["cells"], rows)
        expected = pipeline.compute_piles_candidates([small_atlas[rows]], features=300, random_seed=123456)[0]
        assert np.array_equal(result["communities"], expected[0]) and result["score"] == expected[1]
        assert result["seeds"] == expected[2] and result["k"] == expected[3]


def test_phase_with_the_whole_loop_and_deviants(small_atlas):
    """`candidates="loop", deviants=True`: every pile's candidates == the oracle's loop on the pile's graph, its deviants
    == the reference's compute_cell_gaps driven by the oracle's restatement of the Python around it."""
    from metacells_b200 import divide_and_conquer as dc, piles as pl, pipeline
    from oracle import oracle

    ref = oracle.load_ref()
    if ref is None or not hasattr(ref, "compute_cell_gaps"):
        pytest.skip("oracle/_ref (with gaps.cpp) is not built")
    pile_of_cells = pl.random_piles(2600, 900, random_seed=123456).astype("int32")
    arguments = dict(target_metacell_umis=int(48 * float(small_atlas.sum()) / 2600))
    with dc.Atlas.from_host(small_atlas) as atlas:
        phase = dc.compute_piles_phase(atlas, pile_of_cells, features=300, random_seed=123456, candidates="loop",
                                       deviants=True, candidates_arguments=arguments)
    for pile, result in enumerate(phase):
        rows = np.flatnonzero(pile_of_cells == pile)
        umis = small_atlas[rows]
        graph, cell_umis = pipeline.pile_knn_graph_on_device(umis, features=300, random_seed=123456)
        try:
            outgoing, _ = graph.to_scipy()
        finally:
            graph.close()
        expected = oracle.compute_candidate_metacells(outgoing, cell_umis, random_seed=123456, **arguments)
        assert np.array_equal(result["communities"], expected)
        assert np.array_equal(result["deviants"], oracle.gaps_deviant_cells(umis, expected, ref))


def test_two_phase_walk(small_atlas):
    """Preliminary piles -> metacell groups -> final piles: every cell ends in exactly one final pile and one candidate
    metacell, cells of one preliminary metacell stay together, and the walk is reproducible."""
    from metacells_b200 import divide_and_conquer as dc

    with dc.Atlas.from_host(small_atlas) as atlas:
        first = dc.divide_and_conquer_graphs(atlas, features=300, target_pile_size=650)
        second = dc.divide_and_conquer_graphs(atlas, features=300, target_pile_size=650)
    assert len(first["preliminary"]) == 4
    final = first["final_pile_of_cells"]
    assert final.min() >= 0 and np.array_equal(final, second["final_pile_of_cells"])
    assert np.array_equal(first["metacell_of_cells"], second["metacell_of_cells"]) and first["metacell_of_cells"].min() >= 0
    preliminary_metacells, _ = dc.metacells_of_phase(2600, first["preliminary"])
    for metacell in range(first["preliminary_metacells"]):
        assert len(np.unique(final[preliminary_metacells == metacell])) == 1
    assert sorted(np.concatenate([pile["cells"] for pile in first["final"]]).tolist()) == list(range(2600))
    for key in ("preliminary", "final"):
        assert first["times"][key]["graphs_s"] > 0
    assert first["times"]["groups_s"] > 0


def test_two_phase_walk_with_the_whole_pile_pipeline(small_atlas):
    """`pile_steps="direct"`: every pile goes through the candidates loop, deviants and dissolve; the walk is reproducible,
    the outliers of the preliminary phase are in no final pile, everything else in exactly one."""
    from metacells_b200 import divide_and_conquer as dc

    with dc.Atlas.from_host(small_atlas) as atlas:
        first = dc.divide_and_conquer_graphs(atlas, features=300, target_pile_size=650, pile_steps="direct")
        second = dc.divide_and_conquer_graphs(atlas, features=300, target_pile_size=650, pile_steps="direct")
    assert np.array_equal(first["metacell_of_cells"], second["metacell_of_cells"])
    assert np.array_equal(first["final_pile_of_cells"], second["final_pile_of_cells"])
    preliminary_metacells, count = dc.metacells_of_phase(2600, first["preliminary"])
    assert count == first["preliminary_metacells"] and all("deviants" in pile and "metacells" in pile for pile in first["preliminary"])
    final = first["final_pile_of_cells"]
    assert np.array_equal(final >= 0, preliminary_metacells >= 0)
    for metacell in range(count):
        assert len(np.unique(final[preliminary_metacells == metacell])) == 1
    inside = first["metacell_of_cells"] >= 0
    assert inside.sum() > 0.5 * 2600 and not inside[final < 0].any()


@pytest.mark.parametrize("geo_mean, groups", [(True, True), (True, False), (False, False)])
def test_collect_metacells_matches_oracle(small_atlas, geo_mean, groups):
    """`Atlas.collect_metacells` (pipeline/collect.py:127-232) against the oracle's metacell-by-metacell restatement: cells,
    total UMIs, the zeros' downsampling depth, UMIs per gene and zeros per gene `==`; the fractions (float64 log / exp on the
    device against numpy's) within 1e-6 relative.  The oracle's fractions without the geometric mean are the downsampled
    UMIs' share, which is not provided: only the exact outputs are compared there."""
    from oracle import oracle

    from metacells_b200 import divide_and_conquer as dc

    rng = np.random.default_rng(9)
    metacell_of_cells = rng.integers(0, 70, size=2600).astype("int32")
    metacell_of_cells[rng.random(2600) < 0.05] = -1
    metacell_of_cells[np.flatnonzero(metacell_of_cells == 12)[1:]] = 13  # one of a single cell
    expected = oracle.collect_metacells(small_atlas, metacell_of_cells, random_seed=123456, metacell_geo_mean=geo_mean,
                                        metacell_groups=groups)
    with dc.Atlas.from_host(small_atlas) as atlas:
        mine = atlas.collect_metacells(metacell_of_cells, 70, random_seed=123456, metacell_geo_mean=geo_mean,
                                       metacell_groups=groups, chunk_cells=500)
        whole = atlas.collect_metacells(metacell_of_cells, 70, random_seed=123456, metacell_geo_mean=geo_mean,
                                        metacell_groups=groups)
    filled = expected["grouped"] > 0
    for name in ("grouped", "total_umis"):
        assert np.array_equal(mine[name], expected[name])
    assert np.array_equal(mine["zeros_downsample_umis"][filled], expected["zeros_downsample_umis"][filled])
    assert np.array_equal(mine["umis_per_gene"].cpu().numpy(), expected["umis_per_gene"])
    assert np.array_equal(mine["zeros_per_gene"].cpu().numpy()[filled], expected["zeros_per_gene"][filled])
    for name in ("umis_per_gene", "zeros_per_gene", "fraction_per_gene"):
        assert np.array_equal(mine[name].cpu().numpy(), whole[name].cpu().numpy()), name  # the chunking does not show
    if geo_mean:
        fractions = mine["fraction_per_gene"].cpu().numpy()[filled]
        theirs = expected["fraction_per_gene"][filled]
        assert np.array_equal(fractions == 0, theirs == 0)
        assert np.allclose(fractions, theirs, rtol=1e-6, atol=0)
