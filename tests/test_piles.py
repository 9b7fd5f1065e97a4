"""CPU tests of the pile scheduling that shards the path across GPUs (metacells_b200/piles.py): the reference's
own `test_random_piles` vector, the assignment of piles to ranks, and a world_size-2 `gloo` run of `map_piles`
(the N>1 path of bench.py and of a divide-and-conquer driver; NCCL on the GPU box, gloo here)."""
import os
import socket

import numpy as np
import pytest

from metacells_b200.piles import map_piles, piles_of_rank, random_piles


def test_random_piles_reference_vector():
    # tests/test_utilities.py:468-471 of the reference
    result = random_piles(10, target_pile_size=3, random_seed=123456)
    assert np.array_equal(result, np.array([2, 2, 1, 1, 1, 0, 0, 2, 0, 0]))


@pytest.mark.parametrize("elements,target", [(1, 5), (7, 7), (100, 30), (500_000, 16_000), (1_000_000, 16_000), (23, 2)])
def test_random_piles_sizes(elements, target):
    piles = random_piles(elements, target, random_seed=123456)
    sizes = np.bincount(piles)
    assert sizes.sum() == elements and sizes.min() >= 1
    assert sizes.max() - sizes.min() <= 1                   # computation.py:2311-2318
    ratio = elements / target
    assert len(sizes) in (max(int(np.floor(ratio)), 1), int(np.ceil(ratio)))
    assert np.array_equal(piles, random_piles(elements, target, random_seed=123456))


@pytest.mark.parametrize("world", [1, 2, 3, 8])
def test_piles_of_rank_is_a_partition(world):
    sizes = np.random.default_rng(world).integers(8_000, 24_000, size=63)
    shares = [piles_of_rank(sizes, rank, world) for rank in range(world)]
    assert sorted(sum(shares, [])) == list(range(63))
    loads = [float((sizes[share].astype("float64") ** 2).sum()) for share in shares]
    assert max(loads) <= min(loads) + float(sizes.max()) ** 2   # longest-processing-time-first bound


def _free_port() -> int:
    with socket.socket() as handle:
        handle.bind(("127.0.0.1", 0))
        return handle.getsockname()[1]


def _worker(rank: int, world: int, port: int, queue) -> None:
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        pile_of_cells = random_piles(5_000, 400, random_seed=123456)
        masses = np.arange(5_000, dtype="float64")
        processed = []

        def process(pile, cells):
            processed.append(pile)
            return {"pile": pile, "cells": int(cells.size), "mass": float(masses[cells].sum()), "rank": rank}

        results = map_piles(pile_of_cells, process, rank=rank, world_size=world)
        queue.put((rank, processed, results))
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_map_piles_world_size_two_gloo():
    import torch.multiprocessing as mp

    context = mp.get_context("spawn")
    queue = context.Queue()
    port = _free_port()
    workers = [context.Process(target=_worker, args=(rank, 2, port, queue)) for rank in range(2)]
    for worker in workers:
        worker.start()
    outputs = sorted(queue.get(timeout=120) for _ in workers)
    for worker in workers:
        worker.join(timeout=60)
        assert worker.exitcode == 0
    (_, first_processed, first_results), (_, second_processed, second_results) = outputs
    pile_of_cells = random_piles(5_000, 400, random_seed=123456)
    piles_count = int(pile_of_cells.max()) + 1
    assert sorted(first_processed + second_processed) == list(range(piles_count))   # disjoint, complete
    assert first_processed and second_processed
    assert first_results == second_results                                            # every rank has every result
    alone = map_piles(pile_of_cells, lambda pile, cells: int(cells.size))
    assert [entry["cells"] for entry in first_results] == alone
    assert [entry["rank"] for entry in first_results] == [0 if pile in first_processed else 1 for pile in range(piles_count)]


@pytest.mark.parametrize("rows,world", [(10, 3), (2, 4), (16_000, 8), (100_001, 8), (5, 1)])
def test_row_blocks_cover_the_matrix(rows, world):
    from metacells_b200.tools import row_block_of_rank

    blocks = [row_block_of_rank(rows, rank, world) for rank in range(world)]
    assert blocks[0][0] == 0 and blocks[-1][1] == rows
    assert all(blocks[rank][1] == blocks[rank + 1][0] for rank in range(world - 1))
    widths = [last - first for first, last in blocks]
    assert max(widths) - min(widths) <= 1 and widths[0] == max(widths)


def test_a_failed_pile_is_processed_again():
    """A pile whose step raises McbError (a device-side failure) is run once more; the second failure is raised; errors of
    the arguments are not retried."""
    from metacells_b200 import piles as pl
    from metacells_b200._lib import McbError

    pile_of_cells = np.array([0, 1, 0, 2, 1, 2, 2], dtype="int32")
    calls = []

    def flaky(pile, cells):
        calls.append(pile)
        if pile == 1 and calls.count(1) == 1:
            raise McbError("out of memory")
        return int(cells.sum())

    assert pl.map_piles(pile_of_cells, flaky) == [2, 5, 14]
    assert calls.count(1) == 2 and calls.count(0) == 1

    def broken(pile, cells):
        calls.append(("broken", pile))
        raise McbError("still out of memory")

    with pytest.raises(McbError):
        pl.map_piles(pile_of_cells, broken)
    assert calls.count(("broken", 0)) == 2

    def wrong(pile, cells):
        calls.append(("wrong", pile))
        raise ValueError("bad argument")

    with pytest.raises(ValueError):
        pl.map_piles(pile_of_cells, wrong)
    assert calls.count(("wrong", 0)) == 1
