"""GPU test of the divide-and-conquer walk over a 60 000-cell synthetic atlas (bench.py's generator): its grouping pile has
1 865 metacells and k = n - 1, a graph whose float32 row totals are not exact -- summed left to right the weights were off by
4e-6 and `optimize_partitions` refused the graph (csrc/numpy_sum.cuh).  The walk must run and give every cell
a final pile.  (Named to run late: it holds 1.7 GB of page-locked memory while it runs.)"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_walk_over_a_small_atlas_with_a_complete_grouping_graph():
    import torch

    import bench
    from metacells_b200 import divide_and_conquer as dc

    cells, genes = 60000, 27000
    device = torch.device("cuda", 0)
    data, indices, indptr = bench.synth_atlas_share(cells, genes, bench.SEED, 0, cells, device)
    atlas = dc.Atlas.from_host_shares(data, indices, indptr, genes, rank=0, world_size=1, device=device)
    try:
        walk = dc.divide_and_conquer_graphs(atlas, features="select", random_seed=bench.SEED, workers=2)
    finally:
        atlas.close()
        del data, indices
    assert len(walk["preliminary"]) == 8  # random_piles of 60 000 cells at the target pile size of 8 000
    assert "final" in walk and len(walk["final"]) >= 2
    covered = np.concatenate([pile["cells"] for pile in walk["final"]])
    assert len(covered) == len(np.unique(covered)) and len(covered) <= cells
    assert all("communities" in pile and len(pile["communities"]) == len(pile["cells"]) for pile in walk["final"])
