"""Host-side logic of the divide-and-conquer walk (no GPU): pile sizes, the cells of piles, metacell numbering, and the
exchange of per-pile results between ranks (gloo, world size 2)."""
import multiprocessing as mp
import os

import numpy as np
import pytest


def test_target_pile_size():
    from metacells_b200 import divide_and_conquer as dc

    # pipeline/divide_and_conquer.py:141-199: max(48 * 100, 160000 / mean * 100) clamped to [8000, 24000]
    assert dc.compute_target_pile_size(np.full(10, 6400, dtype="float32")) == 8000
    assert dc.compute_target_pile_size(np.full(10, 1000, dtype="float32")) == 16000
    assert dc.compute_target_pile_size(np.full(10, 100, dtype="float32")) == 24000
    assert dc.compute_target_pile_size(np.full(10, 1000, dtype="float32"), min_target_pile_size=20000, max_target_pile_size=20000) == 20000


def test_rows_of_piles_and_metacell_numbers():
    from metacells_b200 import divide_and_conquer as dc

    pile_of_cells = np.array([1, 0, -1, 1, 0, 2, 1], dtype="int32")
    rows = dc._rows_of_piles(pile_of_cells)
    assert [r.tolist() for r in rows] == [[1, 4], [0, 3, 6], [5]]
    assert dc._rows_of_piles(np.array([-1, -1])) == []
    phase = [{"cells": rows[0], "communities": np.array([1, 0], dtype="int32")},
             {"cells": rows[1], "communities": np.array([0, 0, 1], dtype="int32")},
             {"cells": rows[2], "communities": np.array([0], dtype="int32")}]
    metacells, count = dc.metacells_of_phase(7, phase)
    assert count == 5 and metacells.tolist() == [2, 1, -1, 2, 0, 4, 3]


def _gather_worker(rank, world_size, port, queue):
    import torch.distributed as dist

    from metacells_b200 import divide_and_conquer as dc

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world_size)
    mine = {pile: {"cells": np.arange(pile + 1), "k": pile} for pile in range(5) if pile % world_size == rank}
    merged = dc._gather_dictionaries(mine, world_size)
    queue.put((rank, sorted(merged), [int(merged[pile]["cells"].sum()) for pile in sorted(merged)]))
    dist.destroy_process_group()


@pytest.mark.parametrize("world_size", [2, 8])  # 8: five piles over eight ranks, three ranks hold none (the grouping piles of a walk)
def test_results_of_piles_are_gathered_across_ranks(world_size):
    context = mp.get_context("spawn")
    queue = context.Queue()
    port = 29500 + (os.getpid() + world_size) % 1000
    workers = [context.Process(target=_gather_worker, args=(rank, world_size, port, queue)) for rank in range(world_size)]
    for worker in workers:
        worker.start()
    results = [queue.get(timeout=120) for _ in workers]
    for worker in workers:
        worker.join(timeout=60)
    for _, piles, sums in results:
        assert piles == [0, 1, 2, 3, 4] and sums == [0, 1, 3, 6, 10]


def test_collect_plan_follows_numpy():
    """The vectorised per-metacell vectors of `collect_metacells` == the reference's per-metacell numpy calls
    (pipeline/collect.py:145-161, :207-214): np.quantile + Python's round, int64 totals, scipy's weighted gmean."""
    import scipy.stats as ss

    from metacells_b200 import divide_and_conquer as dc

    rng = np.random.default_rng(4)
    cell_umis = rng.integers(300, 20000, size=3000).astype("float32")
    metacell_of_cells = rng.integers(-1, 90, size=3000).astype("int32")
    metacell_of_cells[metacell_of_cells == 40] = 41
    plan = dc.collect_plan(metacell_of_cells, 90, cell_umis, metacell_groups=True)
    assert sorted(plan["rows"].tolist()) == np.flatnonzero(metacell_of_cells >= 0).tolist()
    for metacell in range(90):
        cells = np.flatnonzero(metacell_of_cells == metacell)
        mine = plan["rows"][plan["starts"][metacell]:plan["starts"][metacell + 1]]
        assert np.array_equal(mine, cells) and plan["grouped"][metacell] == len(cells)
        assert np.array_equal(plan["seed_index"][plan["starts"][metacell]:plan["starts"][metacell + 1]], np.arange(len(cells)))
        if len(cells) == 0:
            continue
        umis = cell_umis[cells].astype("float64")
        assert plan["total_umis"][metacell] == np.sum(umis.astype("int64")) == plan["scale"][metacell]
        assert plan["zeros_downsample_umis"][metacell] == round(np.quantile(umis, 0.1))
        assert (plan["samples"][plan["starts"][metacell]:plan["starts"][metacell + 1]] == plan["zeros_downsample_umis"][metacell]).all()
        regularization = ss.mstats.gmean((1 / 16.0) / umis, weights=np.log2(umis))
        assert np.isclose(np.exp(plan["mean_log_regularization"][metacell]), regularization, rtol=1e-13, atol=0)
