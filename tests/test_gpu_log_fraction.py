"""The log-fraction transform (row a2: pipeline/direct.py:242-249 -> utilities/computation.py:679-741) read back from
the device as the dense [cells, features] operand and compared `==` with numpy's own `x += 1/7; np.log2(x)`."""
import numpy as np
import pytest

from oracle import oracle

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("cells,genes,features,seed", [(1, 40, 7, 1), (37, 300, 300, 2), (500, 2500, 400, 3), (1200, 5000, 1999, 4)])
@pytest.mark.parametrize("downsampled", [True, False])
def test_log_fraction_matches_numpy(cells, genes, features, seed, downsampled):
    from synth import synth_umis

    from metacells_b200 import pipeline

    umis = synth_umis(cells, genes, seed=seed, depth_scale=3.0)
    rng = np.random.default_rng(seed)
    feature_genes = np.sort(rng.choice(genes, features, replace=False)).astype("int32")
    with pipeline.Pile(umis) as pile:
        if downsampled:
            pile.downsample(500, 123456)
            values = pile.downsampled()
        else:
            values = umis
        dense = pile.log_fraction(feature_genes)
        # in another order of the genes too (direct.py slices with a mask, but the entry point takes any distinct list)
        shuffled = rng.permutation(feature_genes).astype("int32")
        dense_shuffled = pile.log_fraction(shuffled)
    assert np.array_equal(dense, oracle.log_fraction(values[:, feature_genes].toarray()))
    assert np.array_equal(dense_shuffled, oracle.log_fraction(values[:, shuffled].toarray()))


def test_log_fraction_of_device_selected_genes():
    from synth import synth_umis

    from metacells_b200 import pipeline

    umis = synth_umis(300, 2000, seed=9)
    with pipeline.Pile(umis) as pile:
        pile.downsample(400, 7)
        genes = pile.select_top_genes(250)
        dense = pile.log_fraction(None)
        down = pile.downsampled()
    assert np.array_equal(dense, oracle.log_fraction(down[:, genes].toarray()))


def test_counts_beyond_the_table_are_an_error():
    """Only the table made by numpy's log2 on the host is bit-exact; a count it does not cover (or a value that is not
    an integer count) must not silently take the device's own log2f."""
    import scipy.sparse as sp

    from metacells_b200 import pipeline
    from metacells_b200._lib import McbError

    umis = sp.csr_matrix(np.array([[0, 3, 70000, 1], [2, 0, 0, 5]], dtype="float32"))
    genes = np.arange(4, dtype="int32")
    with pipeline.Pile(umis) as pile:
        assert np.array_equal(pile.log_fraction(genes), oracle.log_fraction(umis.toarray()))  # table up to the largest count
        with pytest.raises(McbError, match="log table"):
            pile.log_fraction(genes, max_count=100)
        with pytest.raises(McbError, match="log table"):
            pile.knn_graph(genes, k=1, max_count=100)
    fractional = sp.csr_matrix(np.array([[0.5, 3, 1, 1], [2, 0, 0, 5]], dtype="float32"))
    with pipeline.Pile(fractional) as pile:
        with pytest.raises(McbError, match="log table"):
            pile.log_fraction(genes, max_count=10)
