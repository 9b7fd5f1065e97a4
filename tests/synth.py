"""Synthetic UMI matrices for tests and bench (SURVEY.md section 8d generator)."""
import numpy as np
import scipy.sparse as sp


def synth_umis(n_cells: int, n_genes: int, seed: int = 123456, depth_scale: float = 1.0) -> sp.csr_matrix:
    """Cell-type mixture, lognormal gene rates, lognormal depth, Poisson counts -> CSR f32/i32/i32."""
    rng = np.random.default_rng(seed)
    n_types = max(4, n_cells // 400)
    base = np.exp(rng.normal(-3.0, 1.6, size=n_genes))
    fold = rng.normal(0.0, 0.7, size=(n_types, n_genes))
    types = rng.integers(0, n_types, size=n_cells)
    depth = rng.lognormal(0.0, 0.4, size=n_cells) * depth_scale
    blocks = []
    step = 2048
    for start in range(0, n_cells, step):
        stop = min(start + step, n_cells)
        rate = base[None, :] * np.exp(fold[types[start:stop]]) * depth[start:stop, None]
        blocks.append(sp.csr_matrix(rng.poisson(rate).astype("float32")))
    matrix = sp.vstack(blocks, format="csr")
    matrix.sort_indices()
    matrix.indices = matrix.indices.astype("int32")
    matrix.indptr = matrix.indptr.astype("int32")
    return matrix
