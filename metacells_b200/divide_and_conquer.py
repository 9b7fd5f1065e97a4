"""The divide-and-conquer walk over an atlas resident in HBM (pipeline/divide_and_conquer.py:931-1079, 1250-1309).

The reference forces the whole cells x genes matrix to CSR once in the parent process (:1286) and every pile of every
phase is ``ut.slice(adata, obs=pile mask)`` of it, handed to a forked worker (:1267, :1291).  Here the atlas is uploaded
once (``Atlas``), every rank holds it in HBM, and a pile of any phase is a row gather on the device
(``mcb_pile_create_from_atlas``): the host link carries the pile's index list, not its UMIs.  With several ranks each
uploads a share of the rows and the shares are exchanged over NVLink (NCCL broadcasts), so the host link of every rank
carries 1/N of the atlas once.

Phases, as ``_compute_divide_and_conquer_subset`` sequences them (:947-1079):

1. preliminary: ``random_piles`` (:947) -> per pile the k-NN graph and the candidate metacells
   (``choose_seeds`` + ``optimize_partitions`` on the graph as it lies on the device);
2. groups (:1016, ``_compute_metacell_groups`` :1153-1245): the UMIs of every preliminary metacell are summed
   (``collect_metacells`` -> ``mcb_atlas_sum_groups``), the metacells become the "cells" of a grouping pile whose masses
   are the metacells' cell counts and whose target mass is the pile size (:1188-1210), and its communities are the final
   piles;
3. final (:1047): the same per-pile path over the final piles.

What this module does NOT do (outside the hot path): the split / merge rounds of
``compute_candidate_metacells`` beyond the first optimisation, deviant detection and dissolving, the outlier levels
(:1062-1079: they re-run the same phases on the cells left without a metacell), and the rare-gene-module piles.  A
grouping of more metacells than fit one pile is done in random piles of metacells (the reference recurses the whole
divide-and-conquer on them, :1216-1226).
"""
from __future__ import annotations

import ctypes
import time
from typing import Dict, List, Optional, Tuple

import numpy as np
import scipy.sparse as sp

from . import piles as pl
from . import pipeline
from . import tools as tl
from ._lib import check, lib

__all__ = ["Atlas", "compute_target_pile_size", "compute_piles_phase", "group_metacells_into_piles",
           "divide_and_conquer_graphs", "metacells_of_phase"]

# metacells/parameters.py
MIN_TARGET_PILE_SIZE = 8000
MAX_TARGET_PILE_SIZE = 24000
TARGET_METACELLS_IN_PILE = 100
PILES_KNN_K_SIZE_FACTOR = 3
PILES_MIN_SPLIT_SIZE_FACTOR = 1.5
PILES_MAX_MERGE_SIZE_FACTOR = 0.15


def compute_target_pile_size(cell_umis: np.ndarray, *, target_metacell_size: int = pipeline.TARGET_METACELL_SIZE,
                             target_metacell_umis: int = pipeline.TARGET_METACELL_UMIS,
                             min_target_pile_size: int = MIN_TARGET_PILE_SIZE,
                             max_target_pile_size: int = MAX_TARGET_PILE_SIZE,
                             target_metacells_in_pile: int = TARGET_METACELLS_IN_PILE) -> int:
    """pipeline/divide_and_conquer.py:141-199."""
    assert 0 < min_target_pile_size <= max_target_pile_size
    assert cell_umis.dtype == "float32"
    by_cells = target_metacell_size * target_metacells_in_pile
    mean_metacell_size = target_metacell_umis / np.mean(cell_umis)
    by_umis = int(round(mean_metacell_size * target_metacells_in_pile))
    return min(max(by_cells, by_umis, min_target_pile_size), max_target_pile_size)


class Atlas:
    """The cells x genes UMI matrix of a divide-and-conquer run, resident in HBM (``mcb_atlas`` of include/mcb200.h)."""

    def __init__(self, handle: ctypes.c_void_p, shape: Tuple[int, int], keep=()):
        self._handle = handle
        self.shape = shape
        self._keep = keep  # device tensors the atlas borrows
        self._cell_umis: Optional[np.ndarray] = None

    @classmethod
    def from_host(cls, umis: sp.csr_matrix) -> "Atlas":
        """Upload a scipy CSR matrix (float32 values; int32 or int64 structure) in one call."""
        umis = sp.csr_matrix(umis)
        data = np.ascontiguousarray(umis.data, dtype="float32")
        indices = np.ascontiguousarray(umis.indices, dtype="int32")
        indptr = np.ascontiguousarray(umis.indptr, dtype="int64")
        handle = ctypes.c_void_p()
        function = lib().mcb_atlas_create_host
        function.restype = ctypes.c_int
        check(function(ctypes.c_void_p(data.ctypes.data), ctypes.c_void_p(indices.ctypes.data),
                       ctypes.c_void_p(indptr.ctypes.data), ctypes.c_int64(umis.shape[0]), ctypes.c_int64(umis.shape[1]),
                       ctypes.byref(handle)))
        return cls(handle, umis.shape)

    @classmethod
    def from_device(cls, data, indices, indptr_device, indptr_host: np.ndarray, genes: int) -> "Atlas":
        """Borrow CUDA ``torch`` tensors (float32 values, int32 gene indices, int64 row pointers) that already hold the
        atlas; ``indptr_host`` is the same row pointers as a numpy int64 array."""
        indptr_host = np.ascontiguousarray(indptr_host, dtype="int64")
        cells = len(indptr_host) - 1
        handle = ctypes.c_void_p()
        function = lib().mcb_atlas_create_dev
        function.restype = ctypes.c_int
        check(function(ctypes.c_void_p(data.data_ptr()), ctypes.c_void_p(indices.data_ptr()),
                       ctypes.c_void_p(indptr_device.data_ptr()), ctypes.c_void_p(indptr_host.ctypes.data),
                       ctypes.c_int64(cells), ctypes.c_int64(int(genes)), ctypes.byref(handle)))
        return cls(handle, (cells, int(genes)), keep=(data, indices, indptr_device))

    @property
    def handle(self) -> ctypes.c_void_p:
        return self._handle

    @property
    def cell_umis(self) -> np.ndarray:
        """direct.py:214-218: the total UMIs of every cell (float32), summed on the device once."""
        if self._cell_umis is None:
            totals = np.empty(self.shape[0], dtype="float32")
            function = lib().mcb_atlas_cell_totals
            function.restype = ctypes.c_int
            check(function(self._handle, ctypes.c_void_p(totals.ctypes.data)))
            self._cell_umis = totals
        return self._cell_umis

    def pile(self, rows: np.ndarray) -> pipeline.Pile:
        rows = np.asarray(rows, dtype="int64")
        in_range = len(rows) > 0 and rows.min() >= 0 and rows.max() < self.shape[0]
        largest = float(self.cell_umis[rows].max(initial=0)) if in_range else None  # a bad row is reported by the library
        return pipeline.Pile.from_atlas(self._handle, rows, self.shape[1], max_cell_total=largest)

    def sum_groups(self, group_of_cells: np.ndarray, groups: int):
        """Per-group UMI totals, a CUDA float32 ``torch`` tensor [groups, genes] (``mcb_atlas_sum_groups``)."""
        import torch

        group_of_cells = np.ascontiguousarray(group_of_cells, dtype="int32")
        assert len(group_of_cells) == self.shape[0]
        sums = torch.empty((int(groups), self.shape[1]), dtype=torch.float32, device="cuda")
        function = lib().mcb_atlas_sum_groups
        function.restype = ctypes.c_int
        check(function(self._handle, ctypes.c_void_p(group_of_cells.ctypes.data), ctypes.c_int64(int(groups)),
                       ctypes.c_void_p(sums.data_ptr()), ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)))
        return sums

    def collect_metacells(self, metacell_of_cells: np.ndarray, metacells_count: int, *, random_seed: int,
                          metacell_geo_mean: bool = True, metacell_umis_regularization: float = 1 / 16.0,
                          zeros_cell_size_quantile: float = 0.1, metacell_groups: bool = False,
                          chunk_cells: int = 16384) -> dict:
        """``collect_metacells`` (pipeline/collect.py:127-232) over the resident atlas: for every metacell ``grouped``
        (cells), ``total_umis``, ``zeros_downsample_umis`` (host vectors), and CUDA ``torch`` tensors [metacells, genes]:
        ``umis_per_gene`` (float32), ``zeros_per_gene`` (int32) and ``fraction_per_gene`` (float32; with
        ``metacell_groups`` multiplied by the metacell's total UMIs: the "effective UMIs" the grouping of the metacells
        works on, :229-230).  The metacells are walked in chunks of about ``chunk_cells`` cells, each a pile gathered on the
        device (``mcb_pile_collect_metacells``); cells of a negative metacell are left out."""
        import torch

        metacell_of_cells = np.asarray(metacell_of_cells)
        assert len(metacell_of_cells) == self.shape[0]
        plan = collect_plan(metacell_of_cells, metacells_count, self.cell_umis, metacell_umis_regularization=metacell_umis_regularization,
                            zeros_cell_size_quantile=zeros_cell_size_quantile, metacell_groups=metacell_groups)
        genes = self.shape[1]
        fractions = torch.empty((metacells_count, genes), dtype=torch.float32, device="cuda")
        sums = torch.empty((metacells_count, genes), dtype=torch.float32, device="cuda")
        zeros = torch.empty((metacells_count, genes), dtype=torch.int32, device="cuda")
        function = lib().mcb_pile_collect_metacells
        function.restype = ctypes.c_int
        pointer = lambda array: ctypes.c_void_p(array.ctypes.data)  # noqa: E731
        starts = plan["starts"]
        first = 0
        while first < metacells_count:
            last = first + 1
            while last < metacells_count and starts[last + 1] - starts[first] <= chunk_cells:
                last += 1
            cells = slice(starts[first], starts[last])
            metacells = slice(first, last)
            rows = plan["rows"][cells]
            if len(rows):
                per_cell = {name: np.ascontiguousarray(plan[name][cells]) for name in ("weight", "regularization", "umis", "samples",
                                                                                       "seed_index")}
                local = np.ascontiguousarray(plan["metacell"][cells] - first, dtype="int32")
                per_metacell = {name: np.ascontiguousarray(plan[name][metacells]) for name in ("mean_log_regularization", "weight_sum",
                                                                                               "scale", "total_umis_f64", "grouped_i32")}
                with self.pile(rows) as pile:
                    check(function(pile._handle, pointer(local), ctypes.c_int64(last - first), pointer(per_cell["weight"]),
                                   pointer(per_cell["regularization"]), pointer(per_cell["umis"]), pointer(per_cell["samples"]),
                                   pointer(per_cell["seed_index"]), ctypes.c_uint64(int(random_seed)),
                                   ctypes.c_int(int(bool(metacell_geo_mean))), pointer(per_metacell["mean_log_regularization"]),
                                   pointer(per_metacell["weight_sum"]), pointer(per_metacell["scale"]),
                                   pointer(per_metacell["total_umis_f64"]), pointer(per_metacell["grouped_i32"]),
                                   ctypes.c_void_p(fractions[first].data_ptr()), ctypes.c_void_p(sums[first].data_ptr()),
                                   ctypes.c_void_p(zeros[first].data_ptr())))
            else:
                fractions[metacells], sums[metacells], zeros[metacells] = 0, 0, 0
            first = last
        return {"grouped": plan["grouped"], "total_umis": plan["total_umis"], "zeros_downsample_umis": plan["zeros_downsample_umis"],
                "umis_per_gene": sums, "zeros_per_gene": zeros, "fraction_per_gene": fractions}

    def close(self) -> None:
        if getattr(self, "_handle", None):
            lib().mcb_atlas_free(self._handle)
            self._handle = None
        self._keep = ()

    def __del__(self):
        self.close()

    def __enter__(self):
        return self

    def __exit__(self, *_):
        self.close()


def _atlas_from_host_shares(cls, share_data, share_indices, share_indptr: np.ndarray, genes: int, *, rank: int,
                            world_size: int, device=None, times: Optional[dict] = None) -> "Atlas":
    """Every rank holds a contiguous share of the atlas' rows in (page-locked) host memory -- rank r the rows after
    those of rank r - 1 -- uploads it, and the shares are exchanged between the GPUs (one ``torch.distributed``
    broadcast per share and array: NCCL over NVLink), so that every rank ends with the whole atlas in HBM while its
    host link carried only its share.  ``share_data`` / ``share_indices``: float32 / int32 ``torch`` CPU tensors or
    numpy arrays of the share's non-zeros; ``share_indptr``: the share's own row pointers (starting at 0)."""
    import torch
    import torch.distributed as dist

    started = time.perf_counter()
    device = device if device is not None else torch.device("cuda", torch.cuda.current_device())
    share_indptr = np.ascontiguousarray(share_indptr, dtype="int64")
    if world_size > 1:
        every_indptr: List[Optional[np.ndarray]] = [None] * world_size
        dist.all_gather_object(every_indptr, share_indptr)
    else:
        every_indptr = [share_indptr]
    sizes = [np.diff(pointers) for pointers in every_indptr]  # type: ignore[arg-type]
    indptr = np.zeros(sum(len(size) for size in sizes) + 1, dtype="int64")
    np.cumsum(np.concatenate(sizes), out=indptr[1:])
    edges = np.cumsum([0] + [int(pointers[-1]) for pointers in every_indptr])  # type: ignore[index]
    nnz = int(indptr[-1])

    def as_tensor(array, dtype):
        return array if isinstance(array, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(array, dtype=dtype))

    host_data, host_indices = as_tensor(share_data, "float32"), as_tensor(share_indices, "int32")
    device_data = torch.empty(nnz, dtype=torch.float32, device=device)
    device_indices = torch.empty(nnz, dtype=torch.int32, device=device)
    device_indptr = torch.from_numpy(indptr).to(device)
    first, count = int(edges[rank]), int(edges[rank + 1] - edges[rank])
    block = 64 << 20  # elements per copy
    for start in range(0, count, block):
        stop = min(start + block, count)
        device_data[first + start:first + stop].copy_(host_data[start:stop], non_blocking=True)
        device_indices[first + start:first + stop].copy_(host_indices[start:stop], non_blocking=True)
    torch.cuda.synchronize(device)
    uploaded = time.perf_counter()
    if world_size > 1:
        for source in range(world_size):
            if edges[source + 1] > edges[source]:
                dist.broadcast(device_data[int(edges[source]):int(edges[source + 1])], src=source)
                dist.broadcast(device_indices[int(edges[source]):int(edges[source + 1])], src=source)
        torch.cuda.synchronize(device)
    if times is not None:
        times["upload_s"] = uploaded - started
        times["exchange_s"] = time.perf_counter() - uploaded
        times["h2d_bytes"] = count * 8 + len(indptr) * 8
        times["exchanged_bytes"] = (nnz - count) * 8 if world_size > 1 else 0
    return cls.from_device(device_data, device_indices, device_indptr, indptr, genes)


Atlas.from_host_shares = classmethod(_atlas_from_host_shares)  # type: ignore[attr-defined]


def quantile_of_groups(sorted_values: np.ndarray, starts: np.ndarray, quantile: float) -> np.ndarray:
    """``np.quantile(group, quantile)`` (the default linear method) of every group of a float64 vector whose groups lie one
    after the other (``starts``: their offsets, one more than groups), each sorted ascending; NaN for an empty group.
    numpy's own steps (``_quantile``: virtual index ``quantile * (n - 1)``, ``_lerp`` with its switch at one half) on all
    groups at once."""
    sizes = np.diff(starts)
    filled = sizes > 0
    index = quantile * (np.maximum(sizes, 1) - 1)
    previous = np.floor(index).astype("int64")
    gamma = index - previous
    following = np.minimum(previous + 1, np.maximum(sizes, 1) - 1)
    low = sorted_values[np.minimum(starts[:-1] + previous, len(sorted_values) - 1)] if len(sorted_values) else np.zeros(len(sizes))
    high = sorted_values[np.minimum(starts[:-1] + following, len(sorted_values) - 1)] if len(sorted_values) else np.zeros(len(sizes))
    difference = high - low
    result = low + difference * gamma
    late = gamma >= 0.5
    result[late] = (high - difference * (1 - gamma))[late]
    return np.where(filled, result, np.nan)


def collect_plan(metacell_of_cells: np.ndarray, metacells_count: int, cell_umis: np.ndarray, *,
                 metacell_umis_regularization: float = 1 / 16.0, zeros_cell_size_quantile: float = 0.1,
                 metacell_groups: bool = False) -> dict:
    """The per-cell and per-metacell vectors of ``collect_metacells`` that do not touch the genes (pipeline/collect.py:
    127-161, :207-214), for all metacells at once: the cells in metacell order (``rows``, ``metacell``, ``starts``), per cell
    ``umis`` (float64), ``weight`` (log2 UMIs), ``regularization``, ``samples`` (the metacell's ``zeros_downsample_umis``),
    ``seed_index`` (the cell's place in its metacell); per metacell ``grouped``, ``total_umis`` (int64),
    ``zeros_downsample_umis``, ``mean_log_regularization``, ``weight_sum``, ``scale``."""
    metacell_of_cells = np.asarray(metacell_of_cells)
    order = np.argsort(metacell_of_cells, kind="stable")
    order = order[metacell_of_cells[order] >= 0]
    metacell = metacell_of_cells[order].astype("int64")
    grouped = np.bincount(metacell, minlength=metacells_count).astype("int64")
    starts = np.concatenate([[0], np.cumsum(grouped)])
    umis = np.asarray(cell_umis)[order].astype("float64")  # :128: the cells' totals as float64
    total_umis = np.zeros(metacells_count, dtype="int64")
    np.add.at(total_umis, metacell, umis.astype("int64"))  # :153
    by_size = np.lexsort((umis, metacell))
    quantiles = quantile_of_groups(umis[by_size], starts, zeros_cell_size_quantile)
    zeros_downsample_umis = np.rint(np.where(np.isnan(quantiles), 0, quantiles)).astype("int64")  # :159: Python's round
    with np.errstate(divide="ignore", invalid="ignore"):
        weight = np.log2(umis)  # :211
        regularization = metacell_umis_regularization / umis  # :210
        weighted = np.zeros(metacells_count)
        np.add.at(weighted, metacell, weight * np.log(regularization))
        weight_sum = np.zeros(metacells_count)
        np.add.at(weight_sum, metacell, weight)
        mean_log_regularization = np.where(grouped > 0, weighted / np.where(weight_sum == 0, 1, weight_sum), 0.0)
    return {"rows": order.astype("int64"), "metacell": metacell, "starts": starts, "umis": umis, "weight": weight,
            "regularization": regularization, "samples": zeros_downsample_umis[metacell].astype("int32"),
            "seed_index": (np.arange(len(order)) - starts[metacell]).astype("int32"), "grouped": grouped,
            "grouped_i32": grouped.astype("int32"), "total_umis": total_umis, "total_umis_f64": total_umis.astype("float64"),
            "zeros_downsample_umis": zeros_downsample_umis, "mean_log_regularization": mean_log_regularization,
            "weight_sum": weight_sum, "scale": total_umis.astype("float64") if metacell_groups else np.ones(metacells_count)}


def _rows_of_piles(pile_of_cells: np.ndarray) -> List[np.ndarray]:
    """The cells of every pile in ascending order (the order a boolean mask slices them in, :1267)."""
    pile_of_cells = np.asarray(pile_of_cells)
    order = np.argsort(pile_of_cells, kind="stable")
    ordered = pile_of_cells[order]
    first = int(np.searchsorted(ordered, 0))  # cells of pile -1 belong to no pile
    piles_count = int(ordered[-1]) + 1 if len(ordered) > first else 0
    edges = np.searchsorted(ordered, np.arange(piles_count + 1))
    return [order[edges[pile]:edges[pile + 1]] for pile in range(piles_count)]


def _samples_of(totals: np.ndarray) -> int:
    """tools/downsample.py:84-93."""
    return int(round(min(max(tl.DOWNSAMPLE_MIN_SAMPLES, np.quantile(totals, tl.DOWNSAMPLE_MIN_CELL_QUANTILE)),
                         np.quantile(totals, tl.DOWNSAMPLE_MAX_CELL_QUANTILE))))


def _gather_dictionaries(mine: Dict[int, dict], world_size: int) -> Dict[int, dict]:
    if world_size == 1:
        return mine
    import torch.distributed as dist

    everyone: List[Optional[Dict[int, dict]]] = [None] * world_size
    dist.all_gather_object(everyone, mine)
    merged: Dict[int, dict] = {}
    for part in everyone:
        merged.update(part)  # type: ignore[arg-type]
    return merged


def metacells_of_phase(cells_count: int, phase: List[dict]) -> Tuple[np.ndarray, int]:
    """The candidate metacell of every atlas cell, numbered pile by pile (-1: the cell was in no pile), and their count."""
    metacell_of_cells = np.full(cells_count, -1, dtype="int32")
    count = 0
    for pile in phase:
        communities = pile.get("metacells", pile["communities"])  # "metacells": after deviants and dissolve (-1 = outlier)
        metacell_of_cells[pile["cells"]] = np.where(communities >= 0, communities + count, -1)
        count += int(communities.max()) + 1
    return metacell_of_cells, count


def compute_piles_phase(atlas: Atlas, pile_of_cells: np.ndarray, *, rank: int = 0, world_size: int = 1, features=2000,
                        random_seed: int = 123456, candidates=True, workers: int = 2,
                        masses: Optional[np.ndarray] = None, bounds=None, k_arguments: Optional[dict] = None,
                        times: Optional[dict] = None, deviants: bool = False,
                        candidates_arguments: Optional[dict] = None, dissolve: bool = False) -> List[dict]:
    """One phase of the walk (``_compute_metacells_of_piles``, :1250-1309) as far as this package goes: for every pile
    of ``pile_of_cells`` (the pile index of every atlas cell, -1 = in no pile) the k-NN graph (downsample -> feature
    genes -> log-fraction -> similarity -> balanced graph, direct.py:214-283) and, with ``candidates``, ``choose_seeds``
    and the first ``optimize_partitions`` on the graph as it lies on the device.  The piles are dealt to the ranks
    largest first (``piles.piles_of_rank``); a rank's ``workers`` host threads take its piles off one list, each pile on
    its thread's own stream, then all its optimisations run side by side, one SM each.  Every pile gets the same
    ``random_seed`` (:1274-1279).  ``masses``: the mass of every cell for the candidates (default: its UMIs, direct.py
    ``cell_umis``).  ``candidates="loop"``: the whole ``compute_candidate_metacells`` loop instead of its opening
    (``metacells_b200.candidates``, with ``candidates_arguments`` for its parameters in place of ``bounds``); ``deviants``:
    then also ``find_deviant_cells`` on every pile (``metacells_b200.deviants``; up to 32 piles side by side, as far as their
    dense matrices fit 64 GB: 3.5 GB for a 16 000-cell pile).  Returns, on every rank, one dictionary per pile in pile order: ``cells`` (atlas rows),
    ``k``, ``edges`` and, with ``candidates``, ``communities`` (int32 per pile cell), ``score``, ``seeds``, and with
    ``deviants`` the boolean ``deviants`` mask; ``dissolve`` (with ``deviants``): then also ``dissolve_metacells``
    (``metacells_b200.dissolve``): ``metacells`` (int32 per pile cell, -1 = outlier) and the ``dissolved`` mask, which is
    what ``compute_direct_metacells`` leaves of a pile (pipeline/direct.py:311-338)."""
    rows_of_piles = _rows_of_piles(pile_of_cells)
    sizes = [len(rows) for rows in rows_of_piles]
    mine = pl.piles_of_rank(sizes, rank, world_size)
    mine.sort(key=lambda pile: (-sizes[pile], pile))  # a dynamic queue, largest first
    cell_umis = atlas.cell_umis
    masses = cell_umis if masses is None else np.ascontiguousarray(masses, dtype="float32")
    bounds = bounds if bounds is not None else pipeline.metacell_bounds()
    k_arguments = k_arguments or {}
    pool = pipeline.worker_pool(max(1, int(workers)))
    started = time.perf_counter()

    def build(pile: int):
        rows = rows_of_piles[pile]
        samples = _samples_of(cell_umis[rows])
        k = pipeline.knn_k_of_pile(len(rows), masses[rows], **k_arguments)
        with atlas.pile(rows) as handle:
            pipeline._start_selection(handle, features)
            handle.downsample(samples, random_seed)
            return pile, handle.knn_graph(pipeline._feature_genes_of(handle, features), k=min(k, len(rows) - 1))

    results: Dict[int, dict] = {}
    for pile in [pile for pile in mine if sizes[pile] < 3]:  # too small for a graph: one community
        results[pile] = {"cells": rows_of_piles[pile], "k": 0, "edges": 0, "communities": np.zeros(sizes[pile], dtype="int32"),
                         "score": 0.0, "seeds": 1}
    graphs = list(pool.map(pl.retrying(build), [pile for pile in mine if sizes[pile] >= 3]))  # a failed pile is done again
    graphs_seconds = time.perf_counter() - started
    try:
        for pile, graph in graphs:
            results[pile] = {"cells": rows_of_piles[pile], "k": graph.k, "edges": graph.nnz}
        if candidates and graphs:
            def seed_partition(entry):
                pile, graph = entry
                pile_masses = masses[rows_of_piles[pile]]
                wanted = max(pipeline.seeds_count_for(float(sum(pile_masses)), len(pile_masses), bounds), 1)
                return pile, graph.seed_partition(pile_masses, bounds, wanted, random_seed=random_seed)

            def whole_loop(entry):
                from .candidates import compute_candidate_metacells

                pile, graph = entry
                communities = compute_candidate_metacells(graph, masses[rows_of_piles[pile]], random_seed=random_seed,
                                                          **(candidates_arguments or {}))
                return pile, (communities, None, int(communities.max()) + 1)

            side_by_side = pipeline.worker_pool(pipeline._side_by_side(None))
            for pile, (communities, score, seeds) in side_by_side.map(whole_loop if candidates == "loop" else seed_partition,
                                                                      graphs):
                results[pile].update(communities=communities, score=score, seeds=seeds)
            if deviants:
                def deviants_of(entry):
                    from .deviants import find_deviant_cells

                    pile = entry[0]
                    rows = rows_of_piles[pile]
                    with atlas.pile(rows) as handle:
                        mask = find_deviant_cells(handle, results[pile]["communities"], total_umis_per_cell=cell_umis[rows])
                        if not dissolve:
                            return pile, mask, None
                        from .dissolve import dissolve_metacells

                        return pile, mask, dissolve_metacells(handle, results[pile]["communities"], mask,
                                                              np.ascontiguousarray(masses[rows], dtype="float32"))

                # the gaps kernel keeps one warp per candidate busy (a pile: a few dozen warps of a whole GPU), so piles
                # run side by side as far as their dense matrices (8 bytes per cell and gene) fit 64 GB
                largest = max(sizes[pile] for pile, _ in graphs)
                beside = int(min(32, max(workers, 64e9 // max(8 * largest * atlas.shape[1], 1))))
                for pile, mask, dissolved in pipeline.worker_pool(beside).map(deviants_of, graphs):
                    results[pile]["deviants"] = mask
                    if dissolved is not None:
                        results[pile]["metacells"], results[pile]["dissolved"] = dissolved
    finally:
        for _, graph in graphs:
            graph.close()
    if times is not None:
        times["graphs_s"] = times.get("graphs_s", 0.0) + graphs_seconds
        times["candidates_s"] = times.get("candidates_s", 0.0) + time.perf_counter() - started - graphs_seconds
    merged = _gather_dictionaries(results, world_size)
    assert sorted(merged) == list(range(len(rows_of_piles))), "some piles were not processed"
    return [merged[pile] for pile in range(len(rows_of_piles))]


def group_metacells_into_piles(atlas: Atlas, metacell_of_cells: np.ndarray, metacells_count: int, *, target_pile_size: int,
                               features: int = 2000, random_seed: int = 123456, times: Optional[dict] = None,
                               grouping_data: str = "collect") -> np.ndarray:
    """``_compute_metacell_groups`` (:1153-1245): the final pile of every cell.  The metacells are collected on the device
    (``Atlas.collect_metacells`` with ``metacell_groups``: every metacell's geometric-mean gene fractions times its total
    UMIs, pipeline/collect.py:196-230; ``grouping_data="sums"``: their plain UMI sums, the round's first form), the
    metacells are the cells of a grouping pile whose
    masses are their cell counts, with target mass = the pile size, target size ``target_metacells_in_pile``,
    ``piles_knn_k_size_factor`` and the piles' split / merge factors (:1188-1210), and the communities of its candidate
    partition are the final piles.  Every rank computes the same grouping.  More metacells than fit one pile are grouped
    in random piles of metacells, whose graphs are built one after the other and whose optimisations run side by side."""
    import torch

    started = time.perf_counter()
    assert grouping_data in ("collect", "sums")
    if grouping_data == "collect":
        # The effective UMIs are rounded to whole numbers: the per-pile path works on counts (the numpy-made log table, the
        # exact integer gene statistics).  The reference keeps the fractional values of the metacells it does not
        # downsample (those under the samples count) and truncates the others (downsample.cpp:26): under half a UMI per
        # gene and metacell either way.
        sums = torch.round(atlas.collect_metacells(metacell_of_cells, metacells_count, random_seed=random_seed,
                                                   metacell_groups=True)["fraction_per_gene"])
    else:
        sums = atlas.sum_groups(metacell_of_cells, metacells_count)
    metacell_sizes = np.bincount(metacell_of_cells[metacell_of_cells >= 0], minlength=metacells_count).astype("float32")
    metacell_umis = sums.sum(dim=1).cpu().numpy().astype("float32")
    bounds = pipeline.metacell_bounds(target_metacell_size=TARGET_METACELLS_IN_PILE, target_metacell_umis=target_pile_size,
                                      min_split_size_factor=PILES_MIN_SPLIT_SIZE_FACTOR,
                                      max_merge_size_factor=PILES_MAX_MERGE_SIZE_FACTOR)
    k_arguments = dict(target_metacell_size=TARGET_METACELLS_IN_PILE, target_metacell_umis=target_pile_size,
                       knn_k_size_factor=PILES_KNN_K_SIZE_FACTOR)
    if metacells_count > MAX_TARGET_PILE_SIZE:
        pile_of_metacells_first = pl.random_piles(metacells_count, MAX_TARGET_PILE_SIZE // 2, random_seed=random_seed)
    else:
        pile_of_metacells_first = np.zeros(metacells_count, dtype="int32")
    group_of_metacells = np.full(metacells_count, -1, dtype="int32")
    graphs = []
    try:
        for rows in _rows_of_piles(pile_of_metacells_first):
            if len(rows) < 3:  # nothing to group
                graphs.append((rows, None))
                continue
            block = sums if len(rows) == metacells_count else sums[torch.from_numpy(rows).to(sums.device)].contiguous()
            with pipeline.Pile.from_dense_device(block.data_ptr(), len(rows), atlas.shape[1]) as pile:
                pile.set_max_cell_total(float(metacell_umis[rows].max()))
                samples = _samples_of(metacell_umis[rows])
                chosen = features
                if chosen == "select" and len(rows) * samples * samples >= 1 << 48:
                    chosen = 2000  # the per-gene sums of squares of so deep a pile pass the 48 bits the kernel keeps
                pipeline._start_selection(pile, chosen)
                pile.downsample(samples, random_seed)
                k = min(pipeline.knn_k_of_pile(len(rows), metacell_sizes[rows], **k_arguments), len(rows) - 1)
                graphs.append((rows, pile.knn_graph(pipeline._feature_genes_of(pile, chosen), k=k)))
            del block

        def seed_partition(entry):
            rows, graph = entry
            if graph is None:
                return np.zeros(len(rows), dtype="int32")
            masses = metacell_sizes[rows]
            wanted = max(pipeline.seeds_count_for(float(sum(masses)), len(rows), bounds), 1)
            return graph.seed_partition(masses, bounds, wanted, random_seed=random_seed)[0]

        groups = 0
        side_by_side = pipeline.worker_pool(pipeline._side_by_side(None))
        for (rows, _), communities in zip(graphs, side_by_side.map(seed_partition, graphs)):
            group_of_metacells[rows] = communities + groups
            groups += int(communities.max()) + 1
    finally:
        for _, graph in graphs:
            if graph is not None:
                graph.close()
    pile_of_cells = np.full(len(metacell_of_cells), -1, dtype="int32")
    grouped = metacell_of_cells >= 0
    pile_of_cells[grouped] = group_of_metacells[metacell_of_cells[grouped]]
    # a community the first optimisation left far above the largest pile the path is built for is cut into random parts
    # (the reference's split rounds, tools/candidates.py:410-530, bound it by the split factor)
    sizes = np.bincount(pile_of_cells[grouped], minlength=groups)
    for pile in np.flatnonzero(sizes > MAX_TARGET_PILE_SIZE):
        cells = np.flatnonzero(pile_of_cells == pile)
        parts = pl.random_piles(len(cells), target_pile_size, random_seed=random_seed)
        extra = parts > 0
        pile_of_cells[cells[extra]] = groups + parts[extra] - 1
        groups += int(parts.max())
    if times is not None:
        times["groups_s"] = times.get("groups_s", 0.0) + time.perf_counter() - started
    return pile_of_cells


def divide_and_conquer_graphs(atlas: Atlas, *, rank: int = 0, world_size: int = 1, features: int = 2000,
                              random_seed: int = 123456, workers: int = 2, target_pile_size: Optional[int] = None,
                              pile_steps: str = "opening") -> dict:
    """Preliminary phase -> groups -> final phase over a resident atlas (:947-1060).  Returns the piles and candidate
    metacells of both phases and the wall time of every step (``times``; seconds on this rank).  ``pile_steps``: what is
    done to every pile after its k-NN graph -- ``"opening"``: seeds + the first partition optimisation; ``"direct"``: all
    of ``compute_direct_metacells`` (the whole candidates loop, deviant cells, dissolve): the cells a pile leaves as
    outliers are then in no group and no final pile (the reference walks them again in its outlier levels, which are
    not built)."""
    assert pile_steps in ("opening", "direct")
    steps = dict(candidates="loop", deviants=True, dissolve=True) if pile_steps == "direct" else {}
    times: dict = {}
    cells = atlas.shape[0]
    if target_pile_size is None:
        target_pile_size = compute_target_pile_size(atlas.cell_umis)
    preliminary_pile_of_cells = pl.random_piles(cells, target_pile_size, random_seed=random_seed).astype("int32")
    phase_times: dict = {}
    preliminary = compute_piles_phase(atlas, preliminary_pile_of_cells, rank=rank, world_size=world_size, features=features,
                                      random_seed=random_seed, workers=workers, times=phase_times, **steps)
    times["preliminary"] = phase_times
    result = {"target_pile_size": int(target_pile_size), "preliminary_pile_of_cells": preliminary_pile_of_cells,
              "preliminary": preliminary, "times": times}
    if len(preliminary) == 1:  # a single pile is its own final phase (:966-981)
        result["metacell_of_cells"], result["metacells"] = metacells_of_phase(cells, preliminary)
        return result
    metacell_of_cells, metacells_count = metacells_of_phase(cells, preliminary)
    final_pile_of_cells = group_metacells_into_piles(atlas, metacell_of_cells, metacells_count, target_pile_size=target_pile_size,
                                                     features=features, random_seed=random_seed, times=times)
    phase_times = {}
    final = compute_piles_phase(atlas, final_pile_of_cells, rank=rank, world_size=world_size, features=features,
                                random_seed=random_seed, workers=workers, times=phase_times, **steps)
    times["final"] = phase_times
    result.update(preliminary_metacells=metacells_count, final_pile_of_cells=final_pile_of_cells, final=final)
    result["metacell_of_cells"], result["metacells"] = metacells_of_phase(cells, final)
    return result
