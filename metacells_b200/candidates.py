"""``compute_candidate_metacells`` (tools/candidates.py:32-289) on a k-NN graph that stays on the GPU.

The reference's loop is host control flow around four kinds of native work: ``choose_seeds`` (:720-749),
``optimize_partitions`` (:831-880), slicing a community's symmetric sub-graph out of the edge weights (:545-546) and its
minimal cut (:548 -> ``ut.min_cut`` -> ``igraph``).  Here a ``CandidatesSession`` keeps the graph, its incoming layout and
the cell UMIs in HBM (``mcb_candidates_*``); the loop sends the community vector (4 bytes per cell) back and forth, takes
the sub-graphs and minimal cuts of ALL communities of a round from one kernel launch each, and decides on the host what the
reference decides on the host, with the reference's own scalar types (``numpy.float32`` UMI sums divided by Python
integers stay float32: a different precision would move a ``ceil`` or a ``round`` at the boundaries).

``igraph`` is not in this image.  ``mcb_candidates_min_cuts`` is Stoer & Wagner's algorithm with stated tie rules: it finds
a cut of the same (minimal) weight as ``Graph.mincut`` but where several cuts share that weight its choice need not be
igraph's.  A caller with igraph passes ``min_cut=`` (a function of one community's dense symmetric weights returning
``(is_second, strength)``) and gets the reference's own choice.
"""
from __future__ import annotations

import ctypes
from math import ceil, floor, sqrt
from typing import Callable, List, Optional, Sequence, Set, Tuple

import numpy as np

from ._lib import check, lib
from .tools import (COOLDOWN_NODE, COOLDOWN_PASS, MAX_SEED_SIZE_QUANTILE, MIN_SEED_SIZE_QUANTILE, KnnGraph)

__all__ = ["CandidatesSession", "compute_candidate_metacells", "cut_strength", "COOLDOWN_PHASE", "MAX_SPLIT_MIN_CUT_STRENGTH",
           "MIN_CUT_SEED_CELLS"]

# metacells/parameters.py:162-188, :649
TARGET_METACELL_SIZE = 48
TARGET_METACELL_UMIS = 160000
MIN_METACELL_SIZE = 12
MIN_SPLIT_SIZE_FACTOR = 2.0
MAX_MERGE_SIZE_FACTOR = 0.5
MAX_SPLIT_MIN_CUT_STRENGTH = 0.1
MIN_CUT_SEED_CELLS = 7
COOLDOWN_PHASE = 0.75


def _pointer(array: np.ndarray) -> ctypes.c_void_p:
    return ctypes.c_void_p(array.ctypes.data)


def _listing(groups: Sequence[np.ndarray]) -> Tuple[np.ndarray, np.ndarray]:
    """``(members, member_offsets)`` of a list of node-index arrays, as the ``mcb_candidates_*`` calls take them."""
    offsets = np.zeros(len(groups) + 1, dtype="int64")
    if groups:
        np.cumsum([len(group) for group in groups], out=offsets[1:])
    members = np.concatenate(groups).astype("int32", copy=False) if groups else np.zeros(0, dtype="int32")
    return np.ascontiguousarray(members), offsets


class CandidatesSession:
    """What the native calls of one ``compute_candidate_metacells`` share, resident on the GPU (``mcb_candidates``): the
    outgoing and incoming layouts of the edge weights (tools/candidates.py:128-143) and ``cell_umis`` (:145-149).  The
    graph must stay alive while the session is."""

    def __init__(self, graph: KnnGraph, cell_umis: np.ndarray):
        cell_umis = np.ascontiguousarray(cell_umis, dtype="float32")
        self.size = graph.shape[0]
        if len(cell_umis) != self.size:
            raise ValueError("cell_umis must have one entry per node")
        self._graph = graph
        self._handle = ctypes.c_void_p()
        function = lib().mcb_candidates_create
        function.restype = ctypes.c_int
        check(function(graph.handle, _pointer(cell_umis), ctypes.byref(self._handle)))

    def close(self) -> None:
        if getattr(self, "_handle", None):
            function = lib().mcb_candidates_free
            function.restype = None
            function(self._handle)
            self._handle = None

    def __del__(self):
        self.close()

    def __enter__(self):
        return self

    def __exit__(self, *_):
        self.close()

    def choose_seeds(self, seed_of_nodes: np.ndarray, max_seeds_count: int, *, random_seed: int,
                     min_seed_size_quantile: float = MIN_SEED_SIZE_QUANTILE,
                     max_seed_size_quantile: float = MAX_SEED_SIZE_QUANTILE) -> int:
        """``_choose_seeds`` (tools/candidates.py:720-749): fills the negative entries of ``seed_of_nodes`` in place."""
        assert seed_of_nodes.dtype == np.int32 and seed_of_nodes.flags.c_contiguous and len(seed_of_nodes) == self.size
        count = ctypes.c_int64(0)
        function = lib().mcb_candidates_choose_seeds
        function.restype = ctypes.c_int
        check(function(self._handle, ctypes.c_uint64(int(random_seed)), ctypes.c_int64(int(max_seeds_count)),
                       ctypes.c_float(min_seed_size_quantile), ctypes.c_float(max_seed_size_quantile), _pointer(seed_of_nodes),
                       ctypes.byref(count)))
        assert seed_of_nodes.min() == 0  # :748
        return int(count.value)

    def optimize_partitions(self, community_of_nodes: np.ndarray, hot_communities: Sequence[int], bounds: Sequence[float], *,
                            cold_temperature: float, random_seed: int, cooldown_pass: float = COOLDOWN_PASS,
                            cooldown_node: float = COOLDOWN_NODE) -> float:
        """``_optimize_partitions`` (tools/candidates.py:831-880): improves ``community_of_nodes`` in place and returns the
        score.  ``bounds``: low / target / high UMIs, then low / target / high size."""
        assert community_of_nodes.dtype == np.int32 and community_of_nodes.flags.c_contiguous
        hot_mask = np.zeros(max(int(community_of_nodes.max()) + 1, 2), dtype="int8")  # :852-854
        hot_mask[list(hot_communities)] = 1
        score = ctypes.c_double(0.0)
        function = lib().mcb_candidates_optimize_partitions
        function.restype = ctypes.c_int
        check(function(self._handle, ctypes.c_uint32(int(random_seed)), *(ctypes.c_double(float(bound)) for bound in bounds),
                       ctypes.c_double(cooldown_pass), ctypes.c_double(cooldown_node), _pointer(community_of_nodes),
                       _pointer(hot_mask), ctypes.c_int64(len(hot_mask)), ctypes.c_double(cold_temperature), None,
                       ctypes.byref(score)))
        assert community_of_nodes.min() == 0  # :879
        return score.value

    def community_weights(self, groups: Sequence[np.ndarray]) -> List[np.ndarray]:
        """For every group of nodes (ascending), ``edge_weights[group, :][:, group]`` plus its transpose, dense float32
        (tools/candidates.py:545-546)."""
        members, offsets = _listing(groups)
        sizes = np.diff(offsets)
        starts = np.concatenate([[0], np.cumsum(sizes * sizes)])
        dense = np.empty(int(starts[-1]), dtype="float32")
        function = lib().mcb_candidates_community_weights
        function.restype = ctypes.c_int
        check(function(self._handle, _pointer(members), _pointer(offsets), ctypes.c_int64(len(groups)), _pointer(dense)))
        return [dense[starts[at]:starts[at + 1]].reshape(sizes[at], sizes[at]) for at in range(len(groups))]

    def min_cuts(self, groups: Sequence[np.ndarray]) -> List[Tuple[np.ndarray, Optional[float]]]:
        """``ut.min_cut`` (utilities/computation.py:2368-2449) of every group's symmetric sub-graph, all in one launch:
        ``(is_second, strength)`` per group (at least two nodes each)."""
        if not groups:
            return []
        members, offsets = _listing(groups)
        is_second = np.empty(len(members), dtype="int8")
        totals = np.empty((len(groups), 3), dtype="float64")
        function = lib().mcb_candidates_min_cuts
        function.restype = ctypes.c_int
        check(function(self._handle, _pointer(members), _pointer(offsets), ctypes.c_int64(len(groups)), _pointer(is_second),
                       _pointer(totals)))
        results = []
        for at in range(len(groups)):
            side = is_second[offsets[at]:offsets[at + 1]].astype(bool)
            first, second, across = (np.float32(total) for total in totals[at])
            results.append((side, cut_strength(len(side) - int(side.sum()), int(side.sum()), first, second, across)))
        return results


def cut_strength(first_size: int, second_size: int, first_total_weight, second_total_weight, cut_total_weight):
    """The scale-invariant strength of a cut (utilities/computation.py:2428-2449): the mean weight of an edge across it
    over the geometric mean of the mean weights inside either side; ``None`` if a side holds no edge.  The totals are the
    float32 sums the reference accumulates."""
    if cut_total_weight == 0:
        return 0.0
    if not (first_total_weight > 0 and second_total_weight > 0):
        return None
    cut_mean_weight = cut_total_weight / (first_size * second_size)
    first_mean_weight = first_total_weight / ((first_size * (first_size - 1)) / 2)
    second_mean_weight = second_total_weight / ((second_size * (second_size - 1)) / 2)
    return cut_mean_weight / sqrt(first_mean_weight * second_mean_weight)


def _members_of(community_of_nodes: np.ndarray) -> List[np.ndarray]:
    """The nodes of every community ``0 .. max``, ascending (``np.where(community_of_nodes == index)[0]`` for each)."""
    count = int(community_of_nodes.max()) + 1
    order = np.argsort(community_of_nodes, kind="stable")
    sizes = np.bincount(community_of_nodes[community_of_nodes >= 0], minlength=count)
    ends = np.cumsum(sizes) + (len(order) - int(sizes.sum()))  # the unassigned (-1) nodes sort first
    return [order[end - size:end] for size, end in zip(sizes, ends)]


def _seeds_count_for(total_umis, total_size, limits: "_Limits") -> int:
    """tools/candidates.py:293-321: how many seeds make metacells inside the size and UMIs limits, halfway."""
    least_by_umis = int(ceil(total_umis / limits.max_umis))
    most_by_umis = int(ceil(total_umis / limits.min_umis))
    least_by_size = int(ceil(total_size / limits.max_size))
    most_by_size = int(ceil(total_size / limits.min_size))
    if most_by_size < least_by_umis:
        return int(round((most_by_size + least_by_umis) / 2))
    if most_by_umis < least_by_size:
        return int(round((most_by_umis + least_by_size) / 2))
    return int(round((max(least_by_size, least_by_umis) + min(most_by_size, most_by_umis)) / 2))


class _Limits:
    """tools/candidates.py:151-160."""

    def __init__(self, target_metacell_size: int, min_metacell_size: int, target_metacell_umis: int,
                 min_split_size_factor: float, max_merge_size_factor: float):
        self.target_size = target_metacell_size
        self.target_umis = target_metacell_umis
        self.lowest_size = min_metacell_size
        self.min_umis = int(floor(target_metacell_umis * max_merge_size_factor))
        self.max_umis = int(ceil(target_metacell_umis * min_split_size_factor)) - 1
        self.min_size = max(min_metacell_size, int(floor(target_metacell_size * max_merge_size_factor)))
        self.max_size = max(int(ceil(target_metacell_size * min_split_size_factor)) - 1, self.min_size + 1)


def _split_large_communities(community_of_nodes: np.ndarray, node_umis: np.ndarray, limits: _Limits, max_size: float,
                             generator: np.random.RandomState) -> Tuple[int, List[int]]:
    """tools/candidates.py:435-482: a community beyond the UMIs or size limit is dissolved but for one random cell of each
    of its random parts, which seed the communities that replace it.  Returns the number of communities added and the
    communities to re-optimise."""
    groups = _members_of(community_of_nodes)
    next_community = len(groups)
    added, hot = 0, []
    for community, nodes in enumerate(groups):
        size = len(nodes)
        umis = np.sum(node_umis[nodes])  # float32, numpy's pairwise sum of the community's cells in node order
        too_large = (umis > limits.max_umis and size >= 3 * limits.min_size) or (size > max_size and umis >= 3 * limits.min_umis)
        if not too_large:
            continue
        parts = max(2, min(int(round(umis / limits.target_umis)), int(round(size / limits.target_size))))
        while True:  # :457-466
            part_of = generator.randint(parts, size=size)
            if len(np.unique(part_of)) == parts:
                break
        community_of_nodes[nodes] = -1
        for part in range(parts):
            part_nodes = nodes[part_of == part]
            kept = part_nodes[generator.randint(len(part_nodes))]
            if part == 0:
                label = community
            else:
                label = next_community
                next_community += 1
                added += 1
            hot.append(label)
            community_of_nodes[kept] = label
    return added, hot


def _cut_weak_communities(session: CandidatesSession, community_of_nodes: np.ndarray, *, max_split_min_cut_strength: float,
                          min_cut_seed_cells: int, must_complete_cover: bool, atomic_candidates: Set[Tuple[int, ...]],
                          min_cut: Optional[Callable]) -> Tuple[int, int, List[int]]:
    """tools/candidates.py:484-519 and :533-589: every community not yet known to hold together is cut at its weakest;
    a weak enough cut splits it in two, or sheds the smaller side when that is under ``min_cut_seed_cells``.  Returns the
    numbers of communities split and cut and the communities to re-optimise."""
    groups = _members_of(community_of_nodes)
    next_community = len(groups)
    keys = [tuple(nodes.tolist()) for nodes in groups]
    asked = [at for at, key in enumerate(keys) if key not in atomic_candidates and len(key) >= 2]
    if min_cut is None:
        answers = session.min_cuts([groups[at] for at in asked])
    else:
        answers = [min_cut(weights) for weights in session.community_weights([groups[at] for at in asked])]
    answer_of = dict(zip(asked, answers))
    split, cut, hot = 0, 0, []
    for community, nodes in enumerate(groups):
        if keys[community] in atomic_candidates:
            continue
        action = "unchanged"
        if community in answer_of:
            is_second, strength = answer_of[community]
            if strength is not None and not strength > max_split_min_cut_strength:
                second_size = int(np.sum(is_second))
                first_size = len(nodes) - second_size
                small_is_second = not first_size < second_size  # :573-576
                if min(first_size, second_size) >= min_cut_seed_cells:
                    community_of_nodes[nodes[is_second]] = next_community
                    action = "split"
                elif not (must_complete_cover and strength > 0):
                    community_of_nodes[nodes[is_second == small_is_second]] = -1
                    action = "cut"
        if action == "unchanged":
            atomic_candidates.add(keys[community])
            continue
        hot.append(community)
        if action == "split":
            hot.append(next_community)
            next_community += 1
            split += 1
        else:
            cut += 1
    return split, cut, hot


def _reduce_communities(session: CandidatesSession, community_of_nodes: np.ndarray, hot_communities: List[int],
                        node_umis: np.ndarray, limits: _Limits, *, cold_temperature: float, random_seed: int,
                        cooldown_pass: float, cooldown_node: float, cooldown_phase: float, increase_phase: float,
                        min_seed_size_quantile: float, max_seed_size_quantile: float, max_split_min_cut_strength: float,
                        min_cut_seed_cells: int, must_complete_cover: bool, atomic_candidates: Set[Tuple[int, ...]],
                        min_cut: Optional[Callable], trace: Optional[list]) -> Tuple[float, float]:
    """tools/candidates.py:324-407: optimise, then split what is too large (or, when nothing is, cut what is weakly held
    together), re-seed and optimise again, colder every time, until a round changes nothing."""
    generator = np.random.RandomState(random_seed)  # :349 (np.random.seed + np.random.randint: the same stream)
    max_size = limits.max_size
    while True:
        bounds = (limits.min_umis, limits.target_umis, limits.max_umis, limits.min_size, limits.target_size, max_size)
        score = session.optimize_partitions(community_of_nodes, hot_communities, bounds, cold_temperature=cold_temperature,
                                            random_seed=random_seed, cooldown_pass=cooldown_pass, cooldown_node=cooldown_node)
        if trace is not None:
            trace.append(("optimize", score, community_of_nodes.copy()))
        cold_temperature = cold_temperature * (1 - cooldown_phase)
        max_size *= increase_phase

        split, hot_communities = _split_large_communities(community_of_nodes, node_umis, limits, max_size, generator)
        cut = 0
        if split == 0:
            split, cut, hot_communities = _cut_weak_communities(
                session, community_of_nodes, max_split_min_cut_strength=max_split_min_cut_strength,
                min_cut_seed_cells=min_cut_seed_cells, must_complete_cover=must_complete_cover,
                atomic_candidates=atomic_candidates, min_cut=min_cut)
        if trace is not None:
            trace.append(("split", split, cut, community_of_nodes.copy()))
        if split + cut == 0:
            return cold_temperature, score

        max_seeds_count = int(community_of_nodes.max()) + 1
        if cut > 0:  # :395-397: the cells that were shed may seed one more community
            hot_communities.append(max_seeds_count)
            max_seeds_count += 1
        session.choose_seeds(community_of_nodes, max_seeds_count, random_seed=random_seed,
                             min_seed_size_quantile=min_seed_size_quantile, max_seed_size_quantile=max_seed_size_quantile)


def _small_communities(community_of_nodes: np.ndarray, node_umis: np.ndarray, limits: _Limits):
    """tools/candidates.py:592-631: the communities too small to be metacells, how many cells and UMIs they hold."""
    small: Set[int] = set()
    nodes_count, nodes_umis = 0, 0
    for community, nodes in enumerate(_members_of(community_of_nodes)):
        size = len(nodes)
        umis = np.sum(node_umis[nodes])
        if (size < limits.lowest_size or (size < limits.min_size and 2 * umis < limits.max_umis)
                or (umis < limits.min_umis and 2 * size < limits.max_size)):
            small.add(community)
            nodes_count += size
            nodes_umis += umis  # int, then float32
    return small, nodes_count, nodes_umis


def _cancel_communities(community_of_nodes: np.ndarray, cancelled: Set[int]) -> int:
    """tools/candidates.py:634-648: the cells of the cancelled communities lose theirs; the others are renumbered
    densely, in order."""
    count = int(community_of_nodes.max()) + 1
    kept = np.array([community not in cancelled for community in range(count)], dtype=bool)
    new_index = np.where(kept, np.cumsum(kept) - 1, -1).astype("int32")
    assigned = community_of_nodes >= 0
    community_of_nodes[assigned] = new_index[community_of_nodes[assigned]]
    return int(kept.sum())


def compute_candidate_metacells(graph: KnnGraph, cell_umis: np.ndarray, *, target_metacell_size: int = TARGET_METACELL_SIZE,
                                min_metacell_size: int = MIN_METACELL_SIZE, target_metacell_umis: int = TARGET_METACELL_UMIS,
                                min_seed_size_quantile: float = MIN_SEED_SIZE_QUANTILE,
                                max_seed_size_quantile: float = MAX_SEED_SIZE_QUANTILE, cooldown_pass: float = COOLDOWN_PASS,
                                cooldown_node: float = COOLDOWN_NODE, cooldown_phase: float = COOLDOWN_PHASE,
                                increase_phase: float = 1.01, min_split_size_factor: float = MIN_SPLIT_SIZE_FACTOR,
                                max_merge_size_factor: float = MAX_MERGE_SIZE_FACTOR,
                                max_split_min_cut_strength: float = MAX_SPLIT_MIN_CUT_STRENGTH,
                                min_cut_seed_cells: int = MIN_CUT_SEED_CELLS, must_complete_cover: bool = False,
                                random_seed: int, min_cut: Optional[Callable] = None, trace: Optional[list] = None) -> np.ndarray:
    """``compute_candidate_metacells`` (tools/candidates.py:32-289) for a ``KnnGraph`` on the GPU: the ``candidate``
    community of every cell (int32).  The parameters are the reference's (``what`` / ``adata`` become ``graph`` and
    ``cell_umis``); ``min_cut``: see the module's docstring; ``trace``: a list that receives every round's state, for the
    parity tests."""
    assert 0.0 < max_merge_size_factor < min_split_size_factor
    assert target_metacell_size > 0 and target_metacell_umis > 0
    assert 0.0 < cooldown_pass < 1.0 and 0.0 <= cooldown_node <= 1.0 and 0.0 < cooldown_phase <= 1.0
    cell_umis = np.ascontiguousarray(cell_umis, dtype="float32")
    limits = _Limits(target_metacell_size, min_metacell_size, target_metacell_umis, min_split_size_factor, max_merge_size_factor)
    size = graph.shape[0]
    total_umis = sum(cell_umis)  # :162: Python's left-to-right sum, in float32

    with CandidatesSession(graph, cell_umis) as session:
        community_of_nodes = np.full(size, -1, dtype="int32")
        seeds_count = session.choose_seeds(community_of_nodes, _seeds_count_for(total_umis, size, limits), random_seed=random_seed,
                                           min_seed_size_quantile=min_seed_size_quantile,
                                           max_seed_size_quantile=max_seed_size_quantile)
        if trace is not None:
            trace.append(("seeds", seeds_count, community_of_nodes.copy()))
        cold_temperature = 1 - cooldown_pass
        old_score, old_communities, old_small_nodes_count = 1e9, community_of_nodes, size
        atomic_candidates: Set[Tuple[int, ...]] = set()
        hot_communities = list(range(seeds_count))
        while True:
            cold_temperature, score = _reduce_communities(
                session, community_of_nodes, hot_communities, cell_umis, limits, cold_temperature=cold_temperature,
                random_seed=random_seed, cooldown_pass=cooldown_pass, cooldown_node=cooldown_node, cooldown_phase=cooldown_phase,
                increase_phase=increase_phase, min_seed_size_quantile=min_seed_size_quantile,
                max_seed_size_quantile=max_seed_size_quantile, max_split_min_cut_strength=max_split_min_cut_strength,
                min_cut_seed_cells=min_cut_seed_cells, must_complete_cover=must_complete_cover,
                atomic_candidates=atomic_candidates, min_cut=min_cut, trace=trace)
            small, small_nodes_count, small_nodes_umis = _small_communities(community_of_nodes, cell_umis, limits)
            if len(small) < 2:
                break
            if (old_small_nodes_count, old_score) <= (small_nodes_count, score):  # :240-246: no better than before
                community_of_nodes = old_communities
                break
            old_score, old_communities, old_small_nodes_count = score, community_of_nodes.copy(), small_nodes_count
            kept_count = _cancel_communities(community_of_nodes, small)
            small_seeds = min(len(small) - 1, _seeds_count_for(small_nodes_umis, small_nodes_count, limits))
            seeds_count = session.choose_seeds(community_of_nodes, kept_count + small_seeds, random_seed=random_seed,
                                               min_seed_size_quantile=min_seed_size_quantile,
                                               max_seed_size_quantile=max_seed_size_quantile)
            if trace is not None:
                trace.append(("reseed", seeds_count, community_of_nodes.copy()))
            hot_communities = list(range(kept_count, seeds_count))
    assert np.all(community_of_nodes >= 0)
    return community_of_nodes
