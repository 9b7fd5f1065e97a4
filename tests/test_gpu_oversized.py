"""GPU tests of the row-block sharded similarity + per-row top-k of an oversized pile (SURVEY.md 8e;
`tools.compute_obs_obs_similarity_top`, C-ABI `mcb_dev_similarity_block_f32` + `mcb_dev_collect_top_f32`).
One GPU here: the blocks of a 3-rank split are computed one after the other and must reproduce the unsharded
result bit for bit; the NCCL all-gather itself is exercised by scripts/run_oversized_pile.py on 2 GPUs."""
import numpy as np
import pytest

from oracle import oracle

pytestmark = pytest.mark.gpu
SEED = 123456


def log_fraction_like(cells, genes, seed):
    rng = np.random.default_rng(seed)
    types = rng.normal(size=(8, genes)).astype("float32")
    counts = rng.poisson(np.exp(1.0 + 0.6 * types[rng.integers(0, 8, size=cells)])).astype("float32")
    return np.log2(counts + np.float32(1 / 7)).astype("float32")


def test_top_similarity_matches_oracle_within_tolerance():
    from metacells_b200 import tools as tl

    dense = log_fraction_like(700, 300, SEED)
    top = 25
    mine = tl.compute_obs_obs_similarity_top(dense, top)
    assert mine.shape == (700, 700) and np.all(np.diff(mine.indptr) == top)
    expected = oracle.similarity(dense)  # abs_pearson, unit diagonal
    rows = np.repeat(np.arange(700), top)
    assert np.abs(mine.data - expected[rows, mine.indices]).max() <= 1e-5
    assert np.all(mine.data[mine.indices == rows] == 1.0)  # every row keeps itself, similarity exactly 1
    # the kept set is the top set of the oracle's similarity up to the tolerance of the correlation
    cut = -np.sort(-expected, axis=1)[:, top - 1]
    kept = np.zeros((700, 700), dtype=bool)
    kept[rows, mine.indices] = True
    assert np.all(expected[kept] >= np.repeat(cut, top) - 2e-5)
    assert np.all(expected[~kept].reshape(700, -1) <= cut[:, None] + 2e-5)
    for row in range(0, 700, 97):
        assert np.all(np.diff(mine.indices[row * top:(row + 1) * top]) > 0)  # ascending columns (top_per.cpp:35)


def test_row_blocks_reproduce_the_unsharded_result():
    import torch

    from metacells_b200 import tools as tl

    dense = log_fraction_like(1111, 257, SEED + 1)
    top = 40
    whole = tl.compute_obs_obs_similarity_top(dense, top, method="pearson")
    device_rows = torch.from_numpy(dense).cuda()
    pieces = []
    for rank in range(3):
        first, last = tl.row_block_of_rank(1111, rank, 3)
        indices, values = tl.similarity_top_block(device_rows, first, last, top, method="pearson")
        pieces.append((indices.cpu().numpy(), values.cpu().numpy()))
    assert np.array_equal(np.concatenate([piece[0] for piece in pieces]).reshape(-1), whole.indices)
    assert np.array_equal(np.concatenate([piece[1] for piece in pieces]).reshape(-1), whole.data)
