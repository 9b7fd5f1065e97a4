"""Piles across GPUs: the host-side scheduling that replaces the reference's ``ut.parallel_map`` over piles
(pipeline/divide_and_conquer.py:1264-1295) with one process per GPU.

Piles are disjoint sets of cells that are processed without any communication, so the data path has no
collective: every rank takes a share of the piles, runs the per-pile hot path on its own GPU and only the
small per-pile results (a graph handle stays on its GPU; sizes, scores, partitions are gathered) travel, through
``torch.distributed`` (NCCL on the GPU box, gloo in the CPU tests).  Results do not depend on which rank ran which
pile: the reference gives every pile the same ``random_seed`` (divide_and_conquer.py:1274-1279) and so do we.
"""
from __future__ import annotations

from math import ceil, floor
from typing import Any, Callable, Dict, List, Optional, Sequence

import numpy as np

__all__ = ["random_piles", "piles_of_rank", "map_piles", "retrying"]


def random_piles(elements_count: int, target_pile_size: int, *, random_seed: int) -> np.ndarray:
    """``ut.random_piles`` (utilities/computation.py:2270-2325): the pile index of every element, piles of
    roughly ``target_pile_size``; the count of piles is whichever of floor / ceil of the ratio distorts the
    pile size least, and the assignment is a legacy-``np.random`` permutation seeded with ``random_seed``."""
    assert elements_count > 0
    assert target_pile_size > 0
    ratio = elements_count / target_pile_size
    fewer, more = max(floor(ratio), 1), ceil(ratio)
    if fewer == more:
        piles_count = fewer
    else:
        stretch = (elements_count / fewer) / target_pile_size   # piles larger than the target by this factor
        shrink = target_pile_size / (elements_count / more)     # piles smaller than the target by this factor
        assert stretch >= 1 and shrink >= 1
        piles_count = fewer if stretch < shrink else more
    base_size = elements_count // piles_count
    leftover = elements_count - base_size * piles_count
    assert 0 <= leftover < piles_count
    # The reference lists the left-over elements (one for each of the first piles) before the full piles.
    ordered = np.concatenate([np.arange(leftover), np.repeat(np.arange(piles_count), base_size)])
    assert ordered.size == elements_count
    np.random.seed(random_seed)
    return np.random.permutation(ordered)


def piles_of_rank(pile_sizes: Sequence[int], rank: int, world_size: int) -> List[int]:
    """The piles rank ``rank`` of ``world_size`` processes: longest-processing-time-first on the cost model of
    the path (the similarity and rank matrices are quadratic in the pile size), ties by pile index, so that
    every rank computes the same assignment without talking to the others."""
    assert 0 <= rank < world_size
    load = [0.0] * world_size
    mine: List[int] = []
    for pile in sorted(range(len(pile_sizes)), key=lambda index: (-int(pile_sizes[index]), index)):
        target = min(range(world_size), key=lambda index: (load[index], index))
        load[target] += float(pile_sizes[pile]) ** 2
        if target == rank:
            mine.append(pile)
    return sorted(mine)


def retrying(process: Callable[..., Any], attempts: int = 2) -> Callable[..., Any]:
    """``process`` run again (up to ``attempts`` times in all) when it raises an ``McbError``: a pile is a pure function of
    its cells and the seed, so a pile whose step failed on the device (an allocation under memory pressure from the piles
    beside it, a transient launch failure) is simply done again; the last failure is raised.  Argument errors
    (``ValueError``, ``AssertionError``...) are not retried."""
    from ._lib import McbError

    def process_again(*arguments, **keywords):
        for attempt in range(attempts):
            try:
                return process(*arguments, **keywords)
            except McbError:
                if attempt + 1 == attempts:
                    raise
        raise AssertionError("unreachable")

    return process_again


def map_piles(pile_of_cells: np.ndarray, process_pile: Callable[[int, np.ndarray], Any], *, rank: int = 0,
              world_size: int = 1, gather: Optional[Callable[[Dict[int, Any]], List[Dict[int, Any]]]] = None,
              attempts: int = 2) -> List[Any]:
    """Run ``process_pile(pile_index, cell_indices)`` for the piles of this rank and return the results of ALL
    piles in pile order on every rank.  ``gather`` exchanges the per-rank dictionaries; the default uses
    ``torch.distributed.all_gather_object`` on the default process group (no exchange when ``world_size`` is 1).  A pile
    whose processing raises an ``McbError`` is processed again (``attempts`` times in all, see ``retrying``)."""
    process_pile = retrying(process_pile, attempts)
    pile_of_cells = np.asarray(pile_of_cells)
    piles_count = int(pile_of_cells.max()) + 1
    sizes = np.bincount(pile_of_cells, minlength=piles_count)
    results: Dict[int, Any] = {}
    for pile in piles_of_rank(sizes, rank, world_size):
        results[pile] = process_pile(pile, np.flatnonzero(pile_of_cells == pile))
    if world_size > 1:
        if gather is None:
            import torch.distributed as dist

            def gather(mine: Dict[int, Any]) -> List[Dict[int, Any]]:
                everyone: List[Optional[Dict[int, Any]]] = [None] * world_size
                dist.all_gather_object(everyone, mine)
                return everyone  # type: ignore[return-value]

        merged: Dict[int, Any] = {}
        for part in gather(results):
            overlap = merged.keys() & part.keys()
            assert not overlap, f"piles {sorted(overlap)} were processed twice"
            merged.update(part)
        results = merged
    assert sorted(results) == list(range(piles_count)), "some piles were not processed"
    return [results[pile] for pile in range(piles_count)]
