"""The host part of the feature-gene selection (metacells_b200/select.py) against the oracle's restatement of
pipeline/select.py + tools/high.py, with the device statistics replaced by numpy (no GPU)."""
import numpy as np
import pytest
import scipy.sparse as sp

from oracle import oracle


def pandas_window_medians(values, window_size):
    """The reference's own rolling median (computation.py:1885-1890), standing in for the device kernel on the CPU."""
    import pandas as pd

    return pd.Series(values).rolling(window_size).median().values[window_size - 1:]


class HostPile:
    """What `select` asks of a `pipeline.Pile`, computed on the host from a downsampled matrix."""

    def __init__(self, down, top):
        self.down = sp.csc_matrix(down)
        self.shape = down.shape
        self.gene_top_threshold = float(top)

    def gene_statistics(self):
        dense = self.down.toarray().astype("float64")
        totals = dense.sum(axis=0).astype("float32")
        return totals, (dense ** 2).sum(axis=0).astype("uint64"), (dense >= self.gene_top_threshold).sum(axis=0).astype("int32")

    def gene_columns(self, genes):
        values = np.zeros((len(genes), self.shape[0]), dtype="float32")
        sizes = np.zeros(len(genes), dtype="int32")
        for slot, gene in enumerate(genes):
            column = self.down.data[self.down.indptr[gene]:self.down.indptr[gene + 1]]
            values[slot, :len(column)] = column
            sizes[slot] = len(column)
        return values, sizes


@pytest.mark.parametrize("seed", [1, 2])
def test_selection_matches_the_oracle(seed, ref):
    from synth import synth_umis

    from metacells_b200 import select

    umis = synth_umis(900, 2500, seed=seed)
    totals = np.asarray(umis.sum(axis=1)).reshape(-1)
    data = oracle.downsample_compressed(umis.data, umis.indptr, np.full(900, 700, dtype="int32"), 123456)
    down = sp.csr_matrix((data, umis.indices, umis.indptr), shape=umis.shape)
    if seed == 2:  # a gene whose sum of squares passes 2^24
        dense = down.toarray()
        dense[:, 5] = np.random.default_rng(0).integers(150, 260, size=900)
        down = sp.csr_matrix(dense)
    down.eliminate_zeros()
    pile = HostPile(down, 4)
    sums, sums_squared, top_counts = select.gene_sums(pile)
    columns = sp.csc_matrix(down)
    assert np.array_equal(sums, np.asarray(columns.sum(axis=0)).reshape(-1))
    assert np.array_equal(sums_squared, np.asarray(columns.multiply(columns).sum(axis=0)).reshape(-1))
    assert np.array_equal(select.relative_variance_of_genes(sums, sums_squared, 900, window_medians=pandas_window_medians), oracle.relative_variance_per_column(down),
                          equal_nan=True)
    for arguments in ({}, {"min_genes": 1200}, {"min_genes": 2500}, {"min_gene_relative_variance": None, "min_genes": 10},
                      {"min_gene_total": 20, "min_genes": 600}, {"min_gene_top3": None}):
        assert np.array_equal(select.select_feature_genes(pile, window_medians=pandas_window_medians, **arguments), oracle.select_feature_genes(down, ref=ref, **arguments)), arguments
    assert len(totals) == 900
