"""A pile uploaded from its compact host copy (`mcb_pile_create_host_compact`: 16-bit gene indices, 8-bit counts, escapes)
is, on the device, the pile `mcb_pile_create_host` makes: same downsampled values, same k-NN graph, bit for bit."""
import numpy as np
import pytest
import scipy.sparse as sp

from synth import synth_umis

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("cells, genes, chunks", [(1500, 3000, None), (3000, 65536, "5"), (40, 300, None)])
def test_compact_pile_is_the_uploaded_pile(cells, genes, chunks, monkeypatch):
    from metacells_b200 import pipeline

    if chunks:
        monkeypatch.setenv("MCB_PILE_CHUNKS", chunks)
    umis = synth_umis(cells, min(genes, 5000), seed=21, depth_scale=3.0)
    if genes > 5000:  # gene indices up to the 16-bit limit
        columns = np.sort(np.random.default_rng(2).choice(genes, size=umis.shape[1], replace=False))
        columns[-1] = genes - 1
        umis = sp.csr_matrix((umis.data, columns[umis.indices].astype("int32"), umis.indptr), shape=(cells, genes))
    umis = umis.tolil()
    umis[0, 0] = 255          # the first count that does not fit 8 bits
    umis[1, 1] = 254
    umis[2, 3] = 70000
    umis[cells - 1, genes - 1] = 1000
    umis = sp.csr_matrix(umis.tocsr(), dtype="float32")
    umis.sort_indices()
    umis.indices, umis.indptr = umis.indices.astype("int32"), umis.indptr.astype("int32")
    compact = pipeline.compact_csr(umis)
    assert len(compact.escape_at) >= 3 and compact.nbytes < 0.45 * (umis.data.nbytes + umis.indices.nbytes + umis.indptr.nbytes)
    samples, k = 900, min(30, cells - 1)
    with pipeline.Pile(umis) as plain, pipeline.Pile.from_compact(compact) as packed:
        assert packed.shape == plain.shape
        for pile in (plain, packed):
            pile.want_gene_statistics(4.0)
            pile.downsample(samples, 123456)
        expected, mine = plain.downsampled(), packed.downsampled()
        assert np.array_equal(mine.indptr, expected.indptr) and np.array_equal(mine.indices, expected.indices)
        assert np.array_equal(mine.data, expected.data)
        for theirs, ours in zip(plain.gene_statistics(), packed.gene_statistics()):
            assert np.array_equal(theirs, ours)
        genes_chosen = plain.top_total_genes(200)
        graphs = []
        for pile in (plain, packed):
            graph = pile.knn_graph(genes_chosen, k=k)
            try:
                graphs.append(graph.to_scipy()[0])
            finally:
                graph.close()
    assert np.array_equal(graphs[0].indptr, graphs[1].indptr) and np.array_equal(graphs[0].indices, graphs[1].indices)
    assert np.array_equal(graphs[0].data, graphs[1].data)


def test_compact_pile_refusals():
    from metacells_b200 import pipeline
    from metacells_b200._lib import McbError

    wide = sp.csr_matrix((np.ones(1, dtype="float32"), np.array([65536], dtype="int32"), np.array([0, 1], dtype="int32")),
                         shape=(1, 65537))
    with pytest.raises(ValueError, match="65536"):
        pipeline.compact_csr(wide)
    compact = pipeline.compact_csr(synth_umis(50, 200, seed=3))
    compact.escape_at = np.array([5, 2], dtype="int64")
    compact.escape_values = np.array([300, 400], dtype="float32")
    with pytest.raises(McbError, match="sorted"):
        pipeline.Pile.from_compact(compact)
