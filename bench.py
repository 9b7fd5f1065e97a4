#!/usr/bin/env python
"""bench.py -- the per-pile k-NN graph hot path (BASELINE.json metric: k-NN graph cells/sec per pile).

A step = one pass of the hot path over one synthetic pile (SURVEY.md section 8d generator):
downsample -> feature genes -> log-fraction -> abs_pearson similarity (tcgen05) -> rank -> balance ->
prune -> weights.  `value` is measured with the UMI matrix already resident in HBM; `e2e` goes through
the reference-facing host entry points with pinned host buffers (H2D of the CSR pile and D2H of the
graph inside the timed region).  N > 1: every rank builds the graph of its own pile (piles are
independent, divide_and_conquer.py:1264-1283), no data-path collective: weak scaling.

    python bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference ...      the reference's own CPU implementation of the path
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "knn_graph_cells_per_sec_per_pile"
UNIT = "cells/s"
SEED = 123456


def parse_args():
    parser = argparse.ArgumentParser()
    parser.add_argument("--gpus", type=int, default=1)
    parser.add_argument("--steps", type=int, default=10)
    parser.add_argument("--warmup", type=int, default=5)
    parser.add_argument("--impl", default="b200", choices=["b200", "reference"])
    parser.add_argument("--cells", type=int, default=16000)
    parser.add_argument("--genes", type=int, default=27000)
    parser.add_argument("--features", type=lambda text: text if text == "select" else int(text), default="select",
                        help='"select": the feature genes extract_selected_data of pipeline/select.py picks (high top-3, high total, '
                             "high relative variance), in both arms; a number: that many genes of the largest downsampled totals "
                             "(the stand-in of SURVEY.md 8d)")
    parser.add_argument("--cpu-cells", type=int, default=0,
                        help="cells of the pile the CPU legs (--impl reference, cpu_baseline) time; 0 = --cells, the same pile "
                             "as the GPU arm (the path is O(cells^2): any other value is not a like-for-like baseline)")
    parser.add_argument("--no-cpu-baseline", action="store_true")
    parser.add_argument("--atlas-cells", type=int, default=1000000,
                        help="cells of the divide-and-conquer atlas whose piles are timed after the per-pile steps (0: skip)")
    parser.add_argument("--e2e-workers", type=int, default=2,
                        help="host threads of a rank that walk its piles end to end, each on a stream of its own")
    parser.add_argument("--atlas-pile-steps", default="opening", choices=["opening", "direct"],
                        help="what the atlas walk does to a pile after its graph: seeds + first optimisation, or all of "
                             "compute_direct_metacells (candidates loop, deviants, dissolve)")
    parser.add_argument("--atlas-extra", default="",
                        help="comma-separated cell counts of further atlases to walk (BASELINE.json configs C3 / C4: 500000,2000000)")
    parser.add_argument("--oversized-cells", type=int, default=65536,
                        help="N > 1 only: cells of the oversized pile whose similarity is sharded by row blocks over the ranks, "
                             "the per-row top-150 gathered with one NCCL all-gather (0: skip)")
    parser.add_argument("--candidate-piles", type=int, default=16,
                        help="piles whose seeds and first partition optimisation run side by side after the k-NN steps "
                             "(N = 1 only; 0: skip)")
    return parser.parse_args()


def workload_name(args, cells):
    features = ("feature genes selected as pipeline/select.py does" if args.features == "select" else
                f"{args.features} feature genes of the largest totals")
    return (f"pile k-NN graph: {cells} cells x {args.genes} genes synthetic UMIs, {features} "
            f"(downsample+select+log-fraction+correlate+rank+balance+prune_per)")


def config_of(args, cells, genes, k, samples, nnz):
    """The workload description both arms print (the driver compares them)."""
    return {"workload": workload_name(args, cells), "cells": int(cells), "genes": int(genes), "feature_genes": args.features,
            "k": int(k), "downsample_samples": int(samples), "umis_nnz": int(nnz),
            "l2": "inputs larger than L2 (CSR pile 0.45 GB, similarity and rank matrices 1 GB each per step)"}


def graph_digest(weights, indices, indptr):
    """sha256 over the k-NN graph's indptr, indices and float32 weights (tests/test_gpu_bench_pile.py pins the bench
    pile's digest against the oracle and tests/golden/bench_pile.json)."""
    import hashlib

    digest = hashlib.sha256()
    for array in (indptr, indices, weights):
        digest.update(np.ascontiguousarray(array).tobytes())
    return digest.hexdigest()


def golden_digest(cells, genes, features, seed):
    path = os.path.join(ROOT, "tests", "golden", "bench_pile.json")
    if not os.path.exists(path):
        return None
    with open(path) as handle:
        golden = json.load(handle)
    return golden.get(f"{cells}x{genes}x{features}@{seed}")


def make_pile(cells, genes, seed):
    from synth import synth_umis

    from metacells_b200 import pipeline, tools as tl

    umis = synth_umis(cells, genes, seed=seed)
    cell_umis = np.asarray(umis.sum(axis=1)).reshape(-1).astype("float32")
    samples = int(round(min(max(tl.DOWNSAMPLE_MIN_SAMPLES, np.quantile(cell_umis, tl.DOWNSAMPLE_MIN_CELL_QUANTILE)),
                            np.quantile(cell_umis, tl.DOWNSAMPLE_MAX_CELL_QUANTILE))))
    k = pipeline.knn_k_of_pile(cells, cell_umis)
    make_pile.max_cell_total = int(cell_umis.max())  # the pipeline has the totals (samples and k come from them)
    make_pile.samples = samples
    return umis, samples, k


# ---- clocks ------------------------------------------------------------------------------------------


class ClockSampler:
    """nvidia-smi samples during the timed region (B200_PROFILING.md clocks line)."""

    QUERY = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.process = None
        self.lines = []

    def start(self):
        if os.environ.get("MCB_BENCH_NO_CLOCKS"):  # diagnosis only: a run without the sampler is not a bench value
            return
        try:
            self.process = subprocess.Popen(
                ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.process = None

    def _read(self):
        for line in self.process.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.process is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.process.terminate()
        try:
            self.process.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.process.kill()
        clocks, maxima, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.lines:
            parts = [part.strip() for part in line.split(",")]
            if len(parts) < 6:
                continue
            try:
                clocks.append(float(parts[0]))
                maxima.append(float(parts[1]))
            except ValueError:
                continue
            for name, flag in zip(names, parts[2:6]):
                if flag.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(clocks)) if clocks else None, "sm_max_mhz": max(maxima) if maxima else None,
                "samples": len(clocks), "reasons": sorted(reasons)}


# ---- CPU: the reference's own implementation of the path ------------------------------------------------


def cpu_threads():
    try:
        import psutil

        return psutil.cpu_count(logical=False) or os.cpu_count() or 1  # as utilities/parallel.py:127 does
    except ImportError:
        return os.cpu_count() or 1


def cpu_cells_of(args):
    return min(args.cpu_cells, args.cells) if args.cpu_cells > 0 else args.cells


def cpu_step_factory(args):
    """Returns (step, description): step() runs the whole path once on the host cores."""
    from oracle import oracle

    ref = oracle.load_ref()
    threads = cpu_threads() if ref is not None else 1
    if ref is not None:
        ref.set_threads_count(threads)
    cells = cpu_cells_of(args)
    umis, samples, k = make_pile(cells, args.genes, SEED)
    timings = {}

    def step():
        step.graph = oracle.cpu_pile_knn_graph(umis, args.features, k, SEED, ref=ref, timings=timings)
        return cells

    step.config = config_of(args, cells, args.genes, k, samples, umis.nnz)
    info = {"cores": threads, "kind": "reference" if ref is not None else "port",
            "sample": f"one {cells}-cell pile of the same generator ({args.genes} genes, {args.features} feature genes, k={k}), whole path"}
    return step, info, timings


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    step, info, timings = cpu_step_factory(args)
    for _ in range(args.warmup):
        step()
    timings.clear()
    start = time.perf_counter()
    cells = 0
    for _ in range(args.steps):
        cells += step()
    elapsed = time.perf_counter() - start
    value = cells / elapsed
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1000.0 * elapsed / max(args.steps, 1), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "impl": "reference",
        "config": step.config, "timing": "host wall clock", "piles_per_step": 1, "graph_nnz": int(step.graph.nnz),
        "feature_genes_selected": int(timings.get("feature_genes_count", 0)),
        "cpu_baseline": dict(info, value=value, unit=UNIT),
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "stage_seconds_per_step": {name: seconds / max(args.steps, 1) for name, seconds in timings.items()
                                   if name != "feature_genes_count"},
    }
    emit(line)


# ---- GPU arm ---------------------------------------------------------------------------------------------


class Library:
    """ctypes access to the C-ABI entry points the bench drives directly."""

    def __init__(self):
        from metacells_b200 import _lib

        self.lib = _lib.lib()
        self.check = _lib.check
        for name in ("mcb_launch_count", "mcb_graph_nnz", "mcb_graph_k", "mcb_pile_nnz"):
            getattr(self.lib, name).restype = ctypes.c_int64
        self.lib.mcb_launch_count.restype = ctypes.c_longlong

    def call(self, name, *arguments):
        function = getattr(self.lib, name)
        function.restype = ctypes.c_int
        self.check(function(*arguments))

    def profile(self, name):
        total, count = ctypes.c_double(0), ctypes.c_longlong(0)
        self.call("mcb_profile_read", name.encode(), ctypes.byref(total), ctypes.byref(count))
        return total.value, count.value

    def profile_names(self):
        buffer = ctypes.create_string_buffer(4096)
        self.call("mcb_profile_names", buffer, ctypes.c_longlong(4096))
        return [name for name in buffer.value.decode().split("\n") if name]


def p(pointer):
    return ctypes.c_void_p(pointer)


def near_gpu_cpus(local_rank):
    """The host cores next to the rank's GPU (NVML's CPU affinity of the device), or None.  The rank's threads and the
    page-locked buffers they touch first then sit on the GPU's NUMA node: a pile's 0.47 GB upload runs at the link's
    speed instead of crossing the socket interconnect (what `numactl` does for a worker process)."""
    try:
        import pynvml
        import torch

        pynvml.nvmlInit()
        properties = torch.cuda.get_device_properties(local_rank)
        bus = "%08x:%02x:%02x.0" % (properties.pci_domain_id, properties.pci_bus_id, properties.pci_device_id)
        handle = pynvml.nvmlDeviceGetHandleByPciBusId(bus.encode())
        words = pynvml.nvmlDeviceGetCpuAffinity(handle, (os.cpu_count() + 63) // 64)
        cpus = {64 * index + bit for index, word in enumerate(words) for bit in range(64) if word >> bit & 1}
        cpus &= os.sched_getaffinity(0)
        if cpus and cpus != os.sched_getaffinity(0):
            near_gpu_cpus.source = "nvml"
            return cpus
        # NVML names every core (round 1's 8-GPU box did): ask the PCI device for its NUMA node instead
        with open("/sys/bus/pci/devices/%s/numa_node" % bus[4:].lower()) as handle:
            node = int(handle.read())
        if node < 0:
            return cpus or None
        with open("/sys/devices/system/node/node%d/cpulist" % node) as handle:
            listed = set()
            for part in handle.read().strip().split(","):
                first, _, last = part.partition("-")
                listed.update(range(int(first), int(last or first) + 1))
        listed &= os.sched_getaffinity(0)
        near_gpu_cpus.source = "sysfs numa_node %d" % node
        return listed or cpus or None
    except Exception:  # noqa: BLE001 (NVML missing or the attribute absent: stay where the launcher put us)
        return None


near_gpu_cpus.source = "none"


def run_gpu_arm(args):
    import torch
    import torch.distributed as dist

    from metacells_b200 import pipeline, select, tools as tl

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a B200: there is no CPU fallback for the metacells_b200 hot path")
    torch.cuda.set_device(local_rank)
    launcher_cpus = os.sched_getaffinity(0)
    gpu_cpus = None if os.environ.get("MCB_BENCH_NO_AFFINITY") else near_gpu_cpus(local_rank)
    if gpu_cpus:
        os.sched_setaffinity(0, gpu_cpus)  # before the page-locked buffers are allocated and touched
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    library = Library()
    library.call("mcb_set_device", ctypes.c_int(local_rank))
    device = torch.device("cuda", local_rank)

    umis, samples, k = make_pile(args.cells, args.genes, SEED + rank)  # one pile per rank
    max_cell_total = make_pile.max_cell_total
    cells, genes = umis.shape
    table = pipeline.log_fraction_table(samples)
    brf, idf, odf, mod = tl.KNN_BALANCED_RANKS_FACTOR, tl.KNN_INCOMING_DEGREE_FACTOR, tl.KNN_OUTGOING_DEGREE_FACTOR, tl.KNN_MIN_OUTGOING_DEGREE

    # host side: pinned CSR and pinned result buffers (e2e); device side: the resident CSR (value)
    pinned = {name: torch.from_numpy(array).pin_memory() for name, array in
              (("data", umis.data), ("indices", umis.indices), ("indptr", umis.indptr))}
    resident = {name: tensor.to(device) for name, tensor in pinned.items()}
    capacity = cells * (int(round(k * odf)) + 64)
    result = {"weights": torch.empty(capacity, dtype=torch.float32).pin_memory(),
              "indices": torch.empty(capacity, dtype=torch.int32).pin_memory(),
              "indptr": torch.empty(cells + 1, dtype=torch.int32).pin_memory()}
    stream = torch.cuda.current_stream(device)

    select_genes = args.features == "select"
    selected = {"genes": None}  # the feature genes of the last pile (page-lockable, kept alive for the asynchronous copy)

    def hot_path(pile, k=k, meanwhile=None):
        """The path on an existing pile handle; returns the graph handle.  `meanwhile` runs on the host after the
        first (asynchronous) launches: the pipelined legs start the next pile's upload there."""
        library.call("mcb_pile_set_max_cell_total", pile, ctypes.c_uint64(max_cell_total))
        if select_genes:  # the downsample kernel adds up the per-gene statistics of pipeline/select.py
            library.call("mcb_pile_want_gene_statistics", pile, ctypes.c_float(select.SELECT_MIN_GENE_TOP3))
        library.call("mcb_pile_downsample", pile, ctypes.c_int32(samples), ctypes.c_uint64(SEED))
        if select_genes:
            if meanwhile is not None:
                meanwhile()
            view = pipeline.Pile._adopt(pile, (cells, genes))  # the handle stays ours
            view.gene_top_threshold = float(select.SELECT_MIN_GENE_TOP3)
            try:
                chosen = np.flatnonzero(select.select_feature_genes(view)).astype("int32")
            finally:
                view._handle = None
            selected["genes"] = chosen
            selected[threading.get_ident()] = chosen  # one per worker thread
            genes_pointer, genes_count = p(chosen.ctypes.data), len(chosen)
        else:
            # the genes of the largest downsampled totals (the stand-in of SURVEY.md 8d), selected on the device
            library.call("mcb_pile_top_genes", pile, ctypes.c_int64(args.features), None)
            if meanwhile is not None:
                meanwhile()
            genes_pointer, genes_count = None, args.features
        graph = ctypes.c_void_p()
        outside = ctypes.c_int(0)
        library.call("mcb_pile_knn_graph", pile, genes_pointer, ctypes.c_int64(genes_count),
                     p(table.ctypes.data), ctypes.c_int64(len(table)), ctypes.c_float(np.float32(1 / 7)), ctypes.c_int(1),
                     ctypes.c_int64(k), ctypes.c_double(brf), ctypes.c_double(idf), ctypes.c_double(odf), ctypes.c_int64(mod),
                     ctypes.byref(outside), ctypes.byref(graph))
        return graph

    def step_resident(on=None):
        pile = ctypes.c_void_p()
        library.call("mcb_pile_create_dev", p(resident["data"].data_ptr()), p(resident["indices"].data_ptr()),
                     p(resident["indptr"].data_ptr()), ctypes.c_int64(cells), ctypes.c_int64(genes),
                     p((on or stream).cuda_stream), ctypes.byref(pile))
        graph = hot_path(pile)
        nnz = library.lib.mcb_graph_nnz(graph)
        library.lib.mcb_graph_free(graph)
        library.lib.mcb_pile_free(pile)
        return nnz

    def upload():
        """Starts the upload of a pile from the pinned host buffers (row chunks on the pile's copy stream)."""
        pile = ctypes.c_void_p()
        library.call("mcb_pile_create_host", p(pinned["data"].data_ptr()), p(pinned["indices"].data_ptr()),
                     p(pinned["indptr"].data_ptr()), ctypes.c_int64(cells), ctypes.c_int64(genes), ctypes.byref(pile))
        return pile

    # the same pile in the compact host form (16-bit gene indices, 8-bit counts, escapes): 3 bytes per non-zero on the link
    compact = pipeline.compact_csr(umis) if genes <= 65536 else None
    if compact is not None:
        compact_pinned = {name: torch.from_numpy(array).pin_memory() for name, array in
                          (("indices", compact.indices), ("counts", compact.counts),
                           ("escape_at", compact.escape_at if len(compact.escape_at) else np.zeros(1, dtype="int64")),
                           ("escape_values", compact.escape_values if len(compact.escape_at) else np.zeros(1, dtype="float32")))}

    def upload_compact():
        pile = ctypes.c_void_p()
        library.call("mcb_pile_create_host_compact", p(compact_pinned["indices"].data_ptr()), p(compact_pinned["counts"].data_ptr()),
                     p(pinned["indptr"].data_ptr()), ctypes.c_int64(cells), ctypes.c_int64(genes),
                     p(compact_pinned["escape_at"].data_ptr()), p(compact_pinned["escape_values"].data_ptr()),
                     ctypes.c_int64(len(compact.escape_at)), ctypes.byref(pile))
        return pile

    uploader = {"upload": upload}

    def step_e2e(pile=None, meanwhile=None, buffers=None):
        buffers = buffers or result
        if pile is None:
            pile = upload()
        graph = hot_path(pile, meanwhile=meanwhile)
        nnz = library.lib.mcb_graph_nnz(graph)
        if nnz > capacity:
            raise SystemExit(f"result buffers too small: {nnz} > {capacity}")
        library.call("mcb_graph_read", graph, p(buffers["weights"].data_ptr()), None, p(buffers["indices"].data_ptr()),
                     p(buffers["indptr"].data_ptr()))
        library.lib.mcb_graph_free(graph)
        library.lib.mcb_pile_free(pile)
        return nnz

    def barrier():
        torch.cuda.synchronize(device)
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize(device)

    def steps_e2e_overlapped(steps):
        """K end-to-end steps as a worker walks its piles (divide_and_conquer.py:1264-1295 hands piles out one after the
        other): the upload of pile k + 1 is started before the launches of pile k (its buffers are allocated in stream order: behind
        pile k's downsampling the copy would start a millisecond later) and runs under its kernels on the pile's copy
        stream, and the graph of pile k is copied to the host (mcb_graph_read_start, a stream of the graph's own) under the
        first kernels of pile k + 1.  Every step still uploads its pile and reads its graph back inside
        the timed region; only the first upload and the last read are exposed."""
        nnz = 0
        upload = uploader["upload"]
        pending = [upload()]
        reading = []  # the graph whose copy to the host is in flight
        trace = os.environ.get("MCB_BENCH_TRACE")  # developer aid: host times of the stages of every step, on stderr
        sync_read = os.environ.get("MCB_BENCH_SYNC_READ")  # developer aid: read every graph before the next pile starts
        laps = []

        def lap(name):
            if trace:
                laps.append((name, time.perf_counter()))

        def finish_read():
            while reading:
                graph = reading.pop()
                library.call("mcb_graph_read_finish", graph)
                library.lib.mcb_graph_free(graph)

        upload_late = os.environ.get("MCB_BENCH_UPLOAD_AFTER_LAUNCH")  # developer aid: the order measured 0.6 ms slower
        for index in range(steps):
            pile = pending.pop(0)
            lap("start")
            if not upload_late and index + 1 < steps:
                pending.append(upload())

            def meanwhile(last=index + 1 == steps):
                lap("launched")
                if not last and upload_late:
                    pending.append(upload())
                lap("upload issued")
                finish_read()  # before this pile's graph is read into the same host buffers
                lap("previous read")

            graph = hot_path(pile, meanwhile=meanwhile)
            lap("graph")
            nnz = library.lib.mcb_graph_nnz(graph)
            if nnz > capacity:
                raise SystemExit(f"result buffers too small: {nnz} > {capacity}")
            library.call("mcb_graph_read_start", graph, p(result["weights"].data_ptr()), None, p(result["indices"].data_ptr()),
                         p(result["indptr"].data_ptr()))
            reading.append(graph)
            if sync_read:
                finish_read()
            lap("read issued")
            library.lib.mcb_pile_free(pile)
            lap("pile freed")
        finish_read()
        lap("end")
        if trace:
            print("[bench] e2e pipeline: " + " | ".join(f"{name} +{(now - laps[i - 1][1]) * 1e3:.2f}" if i else name
                                                       for i, (name, now) in enumerate(laps)), file=sys.stderr)
        return nnz

    workers = max(1, args.e2e_workers)
    worker_buffers = [result] + [{name: torch.empty_like(tensor).pin_memory() for name, tensor in result.items()}
                                 for _ in range(workers - 1)]

    def steps_e2e_workers(steps):
        """K end-to-end steps by the rank's worker threads: every thread walks its share of the piles one after the other
        (host CSR in, graph out into its own page-locked buffers, as step_e2e) on a stream of its own, so one thread's
        upload, read-back and host-side gaps lie under the other threads' kernels.  Nothing is prefetched and nothing is
        shared between the threads but the GPU (taking turns at the upload was measured: no better here, worse over the
        atlas, where it costs the overlap of a pile's own downsampling with its upload)."""
        import threading

        shares = [steps // workers + (1 if index < steps % workers else 0) for index in range(workers)]
        edges, failures = [0] * workers, []

        def walk(index):
            try:
                library.call("mcb_set_device", ctypes.c_int(local_rank))  # the current device is per thread
                for _ in range(shares[index]):
                    edges[index] = step_e2e(buffers=worker_buffers[index])
            except BaseException as failure:  # noqa: BLE001 (reported by the main thread)
                failures.append(failure)

        threads = [threading.Thread(target=walk, args=(index,)) for index in range(workers)]
        for thread in threads:
            thread.start()
        for thread in threads:
            thread.join()
        if failures:
            raise failures[0]
        return max(edges)

    worker_streams = [torch.cuda.Stream(device) for _ in range(workers)]

    def steps_resident_workers(steps):
        """K resident steps by the rank's worker threads, each on a stream of its own: the host part of one pile's step
        (the numpy tail of the gene selection, 1.7 ms) lies under the kernels of the other thread's pile, as in
        `divide_and_conquer.compute_piles_phase` (workers=2)."""
        import threading

        shares = [steps // workers + (1 if index < steps % workers else 0) for index in range(workers)]
        edges, failures = [0] * workers, []

        def walk(index):
            try:
                library.call("mcb_set_device", ctypes.c_int(local_rank))
                for _ in range(shares[index]):
                    edges[index] = step_resident(on=worker_streams[index])
                worker_streams[index].synchronize()
            except BaseException as failure:  # noqa: BLE001 (reported by the main thread)
                failures.append(failure)

        threads = [threading.Thread(target=walk, args=(index,)) for index in range(workers)]
        for thread in threads:
            thread.start()
        for thread in threads:
            thread.join()
        if failures:
            raise failures[0]
        return max(edges)

    def timed(step, steps, whole_run=False):
        """K steps between events on the launching stream; returns (max-over-ranks ms, nnz)."""
        begin, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        begin.record(stream)
        nnz = 0
        if whole_run:
            nnz = step(steps)
        else:
            for _ in range(steps):
                nnz = step()
        end.record(stream)
        barrier()
        elapsed = torch.tensor([begin.elapsed_time(end)], dtype=torch.float64, device=device)
        print(f"[bench] rank {rank}: {float(elapsed.item()) / steps:.3f} ms per step ({step.__name__})", file=sys.stderr)
        if world > 1:
            dist.all_reduce(elapsed, op=dist.ReduceOp.MAX)
        return float(elapsed.item()), nnz

    for _ in range(max(args.warmup, 3)):
        step_resident()
    step_e2e()
    library.call("mcb_profile_reset")
    library.call("mcb_profile_enable", ctypes.c_int(1))
    sampler = ClockSampler(local_rank)
    sampler.start()
    launches_before = library.lib.mcb_launch_count()
    resident_ms, nnz = timed(step_resident, args.steps)
    launches = library.lib.mcb_launch_count() - launches_before
    clocks = sampler.stop()
    kernel_ms = {name: library.profile(name) for name in library.profile_names()}
    library.call("mcb_profile_enable", ctypes.c_int(0))
    library.call("mcb_profile_reset")
    steps_resident_workers(2 * workers)
    resident_workers_ms, _ = timed(steps_resident_workers, args.steps, whole_run=True)
    for _ in range(2):  # the sampler's exit and the profiler's teardown stay out of the timed region
        step_e2e()
    e2e_sequential_ms, _ = timed(step_e2e, args.steps)
    steps_e2e_overlapped(3)  # two piles in flight: lets the memory pool grow before the timed region
    if os.environ.get("MCB_BENCH_PROFILE_E2E"):  # diagnosis only: kernel times while the next pile's upload is in flight
        library.call("mcb_profile_enable", ctypes.c_int(1))
    e2e_ms, _ = timed(steps_e2e_overlapped, args.steps, whole_run=True)
    if os.environ.get("MCB_BENCH_PROFILE_E2E"):
        print("[bench] kernels under upload:", {name: round(library.profile(name)[0] / args.steps, 3) for name in library.profile_names()},
              file=sys.stderr)
        library.call("mcb_profile_enable", ctypes.c_int(0))
    steps_e2e_workers(2 * workers)
    e2e_workers_ms, _ = timed(steps_e2e_workers, args.steps, whole_run=True)
    e2e_compact = None
    if compact is not None:  # the one-thread prefetch walk again, from the compact host copy
        uploader["upload"] = upload_compact
        steps_e2e_overlapped(3)
        compact_ms, compact_nnz = timed(steps_e2e_overlapped, args.steps, whole_run=True)
        compact_check = step_e2e(pile=upload_compact())
        compact_digest = graph_digest(result["weights"].numpy()[:compact_check], result["indices"].numpy()[:compact_check],
                                      result["indptr"].numpy())
        uploader["upload"] = upload
        e2e_compact = {"ms_per_step": compact_ms / args.steps, "value": world * cells * args.steps / (compact_ms * 1e-3),
                       "h2d_bytes_per_step": int(compact.nbytes), "escapes": int(len(compact.escape_at)),
                       "graph_sha256": compact_digest,
                       "note": "mcb_pile_create_host_compact: the pile kept on the host as 16-bit gene indices + 8-bit counts "
                               "(+ escapes), widened on the device; same one-thread prefetch walk, same graph; not the "
                               "reference's host format (scipy CSR float32 / int32), so it is reported beside the headline"}
    e2e_prefetch_ms = e2e_ms
    # the headline is fixed beforehand (not the faster of the two): one host thread with the next upload in flight
    e2e_mode = "one host thread, next upload and previous read under the current pile's kernels"
    # the graph of one more end-to-end step, as it arrived in the page-locked result buffers, against the digest the
    # test suite pins (tests/test_gpu_bench_pile.py: == oracle on the same pile, == tests/golden/bench_pile.json)
    check_nnz = step_e2e()
    digest = graph_digest(result["weights"].numpy()[:check_nnz], result["indices"].numpy()[:check_nnz], result["indptr"].numpy())
    expected_digest = golden_digest(cells, genes, args.features, SEED + rank)
    graph_check = {"sha256": digest, "expected": expected_digest,
                   "ok": (digest == expected_digest) if expected_digest else None,
                   "source": "tests/golden/bench_pile.json (pinned by tests/test_gpu_bench_pile.py against the oracle)"}
    if expected_digest and digest != expected_digest:
        raise SystemExit(f"bench: the timed pile's k-NN graph differs from the pinned digest ({digest} != {expected_digest})")

    def extra_leg(name, leg):
        """The legs beside the headline (the atlas walk, the oversized pile, the candidates) must not cost the line that is
        already measured: a failure every rank sees alike (a limit, out of memory) is reported under the leg's key.  A
        failure of one rank alone still leaves the others waiting in the leg's next collective."""
        try:
            return leg()
        except Exception as error:  # noqa: BLE001
            print(f"[bench] rank {rank}: {name} failed: {error!r}", file=sys.stderr, flush=True)
            return {"error": f"{type(error).__name__}: {error}"[:400]}

    atlas = None
    if args.atlas_cells > 0:
        resident.clear()  # the resident pile's arrays are not needed any more
        torch.cuda.empty_cache()
        atlas = extra_leg("atlas walk", lambda: time_atlas_walk(args, args.atlas_cells, device, rank, world, barrier))
        extra = [int(text) for text in args.atlas_extra.split(",") if text.strip()]
        if extra and "error" not in atlas:
            atlas["other_sizes"] = [extra_leg("atlas walk", lambda size=size: time_atlas_walk(args, size, device, rank, world, barrier))
                                    for size in extra]

    oversized = None
    if world > 1 and args.oversized_cells > 0:
        oversized = extra_leg("oversized pile", lambda: time_oversized(args, rank, world, device, barrier))

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    peaks = {}
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        with open(peaks_path) as handle:
            peaks = json.load(handle)
    # FP32-emulated tensor peak := dense TF32 peak / 3 (3 MMAs per algorithmic FMA).  The TF32 peak is measured here
    # (torch.matmul with TF32 allowed, 8192^3: cuBLAS as the yardstick only, never on the path).  The contraction runs
    # ~2 ms at a time inside a timed region of a fraction of a second with the clocks at their maximum, so the BURST
    # figure is the denominator (B200_PROFILING.md: burst for a kernel timed alone / in a short region); the fraction
    # against the sustained figure and the tensor-pipe utilisation of the ncu capture are given beside it.
    tf32 = measure_tf32_peak(device)
    peak = tf32["burst_tflops"] / 3.0
    peak_source = "measured in this run: torch.matmul TF32 8192^3 best of 10 (burst) / 3 (3xTF32)"
    hbm = peaks.get("hbm_gbs")
    hbm_source = "MEASURED_PEAKS.json hbm_gbs"
    if hbm is None:
        hbm, hbm_source = 6500.0, "fallback 6.5 TB/s measured copy bandwidth (B200_PROFILING.md)"

    def per_launch_ms(*names):
        """Average duration of one pass of the named kernels (they run once per pile each)."""
        total = 0.0
        for name in names:
            milliseconds, count = kernel_ms.get(name, (0.0, 0))
            if not count:
                return None
            total += milliseconds / count
        return total

    def hbm_line(kernel, names, algorithmic_bytes, traffic, traffic_source):
        ms = per_launch_ms(*names)
        achieved_gbs = algorithmic_bytes / (ms * 1e-3) / 1e9 if ms else None
        return {"bound": "hbm", "kernel": kernel, "achieved": achieved_gbs, "peak": hbm, "unit": "GB/s",
                "frac": achieved_gbs / hbm if achieved_gbs else None, "traffic": traffic, "traffic_source": traffic_source,
                "peak_source": hbm_source, "algorithmic_bytes_per_launch": algorithmic_bytes, "ms_per_launch": ms}

    umma_ms = per_launch_ms("correlate_umma")
    features_count = len(selected["genes"]) if select_genes else args.features
    flops = float(cells) * cells * features_count  # algorithmic: 2 * g * N (N - 1) / 2 ~= N^2 g
    achieved = flops / (umma_ms * 1e-3) / 1e12 if umma_ms else None
    correlate_line = {"bound": "tensor", "kernel": "correlate_umma_kernel", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                      "frac": achieved / peak if achieved else None, "traffic": ncu_traffic("correlate_umma", cells, features_count),
                      "traffic_source": "ncu --set full dram__bytes_read.sum + dram__bytes_write.sum, profiles/r02_ncu_full_correlate_umma.txt (890 genes) / r01_ncu_full_correlate_umma.txt (2 000)",
                      "peak_source": peak_source, "algorithmic_flops_per_launch": flops, "ms_per_launch": umma_ms,
                      "tf32_peak_measured": tf32,
                      "frac_of_sustained": achieved / (tf32["sustained_tflops"] / 3.0) if achieved else None,
                      "frac_of_bf16_burst_over_6": achieved / (peaks["bf16_tflops"] / 6.0) if achieved and peaks.get("bf16_tflops") else None,
                      "ncu_tensor_pipe_pct": {890: 63.4, 2000: 67.0}.get(features_count),
                      "ncu_tensor_pipe_source": "tensor pipe, % of active cycles: profiles/r02_ncu_full_correlate_umma.txt (890 genes), "
                                                "profiles/r01_ncu_full_correlate_umma.txt (2 000 genes)"}
    # SURVEY 8d's algorithmic bytes: ranks read and write the whole square matrix; the downsample reads and writes the
    # pile's values and reads the row pointers; balance reads every rank once (count and fill passes are one stage).
    rank_line = hbm_line("rank_matrix_bucket_kernel", ("rank_matrix",), 8.0 * cells * cells, ncu_traffic("rank_matrix", cells, 2000),
                         "ncu --set full dram__bytes_read.sum + dram__bytes_write.sum, profiles/r02_ncu_full_rank_matrix_bucket.txt")
    downsample_line = hbm_line("downsample_warp_kernel", ("downsample",), 8.0 * umis.nnz + 8.0 * cells + 4, ncu_traffic("downsample", cells, 0),
                               "ncu --set full dram__bytes_read.sum + dram__bytes_write.sum, profiles/r02_ncu_full_downsample_warp.txt")
    balance_line = hbm_line("balance_kernel<count> + balance_scatter_kernel", ("balance_count", "balance_fill"), 4.0 * cells * cells,
                            ncu_traffic("balance", cells, 0),
                            "ncu --set full, count pass (reads the ranks once, parks masks + values) + scatter, "
                            "profiles/r02b_ncu_table_step.md")
    # the dominant kernel is the one with the largest share of the step
    candidates = [line for line in (rank_line, correlate_line, downsample_line, balance_line) if line["ms_per_launch"]]
    dominant = max(candidates, key=lambda line: line["ms_per_launch"]) if candidates else correlate_line
    line = {
        "metric": METRIC, "value": world * cells * args.steps / (resident_ms * 1e-3), "unit": UNIT, "n_gpus": world,
        "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": resident_ms / args.steps, "higher_is_better": True,
        "resident_worker_threads": {"threads": workers, "ms_per_step": resident_workers_ms / args.steps,
                                    "value": world * cells * args.steps / (resident_workers_ms * 1e-3),
                                    "note": "the same K resident steps dealt over the rank's worker threads, one stream each "
                                            "(a pile's host-side gene selection under the other pile's kernels, as the "
                                            "divide-and-conquer walk runs its piles); reported beside `value`, which stays "
                                            "one pile at a time"},
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config_of(args, cells, genes, k, samples, umis.nnz), "graph_nnz": int(nnz), "piles_per_step": world,
        "feature_genes_selected": int(features_count),
        "clocks": clocks,
        "host": {"cpus_near_gpu": len(gpu_cpus) if gpu_cpus else None, "cpus": len(launcher_cpus),
                 "affinity_source": near_gpu_cpus.source},
        "e2e": {"value": world * cells * args.steps / (e2e_ms * 1e-3), "unit": UNIT, "ms_per_step": e2e_ms / args.steps,
                "unpipelined_ms_per_step": e2e_sequential_ms / args.steps,
                "mode": e2e_mode, "graph_check": graph_check, "one_thread_prefetch_ms_per_step": e2e_prefetch_ms / args.steps,
                "worker_threads_ms_per_step": e2e_workers_ms / args.steps, "worker_threads": workers,
                "compact_host_format": e2e_compact,
                "note": "every step uploads its pile from pinned host memory and reads its graph back into pinned memory, both "
                        "inside the timed region, walked as a worker of the divide-and-conquer loop walks its piles, two ways (`value` "
                        "is always the first, see `mode`): one_thread_prefetch_ms_per_step = one host thread, the upload of the next "
                        "pile (chunks on a copy stream) and the read of the previous graph (mcb_graph_read_start) under the "
                        "kernels of the current pile; worker_threads_ms_per_step = `worker_threads` host threads, each pile on its "
                        "thread's own stream, nothing prefetched.  unpipelined_ms_per_step: one pile at a time",
                "h2d_bytes_per_step": int(umis.data.nbytes + umis.indices.nbytes + umis.indptr.nbytes + table.nbytes + 4 * features_count),
                "d2h_bytes_per_step": int(nnz * 8 + (cells + 1) * 4 + genes * (12 if select_genes else 4))},
        "gpu_launches": int(launches),
        "roofline": dominant,
        "roofline_kernels": [line for line in candidates if line is not dominant],
        "kernel_ms_per_step": {name: total / args.steps for name, (total, _) in sorted(kernel_ms.items())},
    }
    if atlas is not None:
        line["atlas"] = atlas
    if oversized is not None:
        line["oversized"] = oversized
    if world == 1 and args.candidate_piles > 0:
        line["candidates"] = extra_leg("candidates", lambda: time_candidates(args, umis))
    os.sched_setaffinity(0, launcher_cpus)  # the CPU baseline uses every core the launcher gave us
    if world == 1 and not args.no_cpu_baseline:
        step, info, _ = cpu_step_factory(args)
        start = time.perf_counter()
        cpu_cells = step()
        line["cpu_baseline"] = dict(info, value=cpu_cells / (time.perf_counter() - start), unit=UNIT)
    emit(line)
    if world > 1:
        dist.destroy_process_group()


def synth_atlas_share(cells, genes, seed, first, last, device, chunk=8192):
    """Rows [first, last) of a synthetic atlas of `cells` distinct cells (the model of tests/synth.py: cell types with
    log-normal gene rates, log-normal depth, Poisson counts), generated block by block on the GPU with torch and moved to
    page-locked HOST memory as CSR float32 / int32 (+ the share's own int64 row pointers): the walk uploads it from there.
    A block depends on (seed, block index) only, so every rank generates its own share of the same atlas."""
    import torch

    generator = torch.Generator(device=device)
    generator.manual_seed(seed)
    types_count = max(4, cells // 400)
    base = torch.exp(torch.randn(genes, generator=generator, device=device) * 1.6 - 3.0)
    fold = torch.randn((types_count, genes), generator=generator, device=device) * 0.7
    on_device, flushed, sizes, total = [], [], [], 0

    def flush():
        if on_device:
            flushed.append((torch.cat([part[0] for part in on_device]).cpu(), torch.cat([part[1] for part in on_device]).cpu()))
            on_device.clear()

    for start in range(first - first % chunk, last, chunk):
        stop = min(start + chunk, cells)
        generator.manual_seed(seed * 1000003 + start)
        types = torch.randint(0, types_count, (stop - start,), generator=generator, device=device)
        depth = torch.exp(torch.randn(stop - start, generator=generator, device=device) * 0.4)
        counts = torch.poisson(base[None, :] * torch.exp(fold[types]) * depth[:, None], generator=generator)
        counts = counts[max(first - start, 0):min(last, stop) - start]
        nonzero = counts.nonzero(as_tuple=True)
        on_device.append((counts[nonzero].to(torch.float32), nonzero[1].to(torch.int32)))
        sizes.append((counts != 0).sum(dim=1).cpu())
        total += int(on_device[-1][0].numel())
        del counts, nonzero
        if len(on_device) >= 8:
            flush()
    flush()
    data = torch.empty(total, dtype=torch.float32).pin_memory()
    indices = torch.empty(total, dtype=torch.int32).pin_memory()
    at = 0
    for values, columns in flushed:
        data[at:at + values.numel()] = values
        indices[at:at + values.numel()] = columns
        at += values.numel()
    indptr = np.zeros(last - first + 1, dtype="int64")
    np.cumsum(torch.cat(sizes).numpy(), out=indptr[1:])
    assert at == total == int(indptr[-1])
    return data, indices, indptr


def time_atlas_walk(args, cells, device, rank, world, barrier):
    """BASELINE.json's second metric: the wall time of a divide-and-conquer walk over a synthetic atlas of `cells` DISTINCT
    cells (configs C3 / C4 and the 1M-cell atlas) at `world` GPUs, strong scaling: every rank's share of the atlas is
    uploaded from page-locked host memory once and exchanged over NVLink (every rank then holds the atlas in HBM), then
    preliminary piles (random_piles) -> k-NN graph + seeds + first partition optimisation of every pile -> the preliminary
    metacells are summed and grouped into the final piles -> the same per-pile path over the final piles.  Piles are row
    gathers on the device; graphs are left on the device and only the candidate partitions are read back."""
    import torch
    import torch.distributed as dist

    from metacells_b200 import divide_and_conquer as dc

    first, last = cells * rank // world, cells * (rank + 1) // world
    started = time.perf_counter()
    data, indices, indptr = synth_atlas_share(cells, args.genes, SEED, first, last, device)
    generate_s = time.perf_counter() - started
    torch.cuda.empty_cache()
    barrier()
    started = time.perf_counter()
    upload = {}
    atlas = dc.Atlas.from_host_shares(data, indices, indptr, args.genes, rank=rank, world_size=world, device=device, times=upload)
    try:
        atlas.cell_umis  # the cells' totals, on the device (direct.py:214-218)
        barrier()
        resident_s = time.perf_counter() - started
        walk = dc.divide_and_conquer_graphs(atlas, rank=rank, world_size=world, features=args.features, random_seed=SEED,
                                            workers=max(1, args.e2e_workers), pile_steps=args.atlas_pile_steps)
        barrier()
        wall_s = time.perf_counter() - started
    finally:
        atlas.close()
        del data, indices
    times = walk["times"]
    flat = [resident_s, wall_s, times["preliminary"]["graphs_s"], times["preliminary"]["candidates_s"], times.get("groups_s", 0.0),
            times.get("final", {}).get("graphs_s", 0.0), times.get("final", {}).get("candidates_s", 0.0), upload["upload_s"],
            upload["exchange_s"]]
    numbers = torch.tensor(flat, dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(numbers, op=dist.ReduceOp.MAX)
    flat = [float(value) for value in numbers.tolist()]
    final = walk.get("final", walk["preliminary"])
    sizes = [len(pile["cells"]) for pile in final]
    return {"metric": "divide_and_conquer_wall_time", "cells": int(cells), "genes": int(args.genes), "n_gpus": world,
            "scaling": "strong", "phases": 2 if "final" in walk else 1, "wall_s": flat[1], "cells_per_s": cells / flat[1],
            "atlas_resident_s": flat[0], "upload_s": flat[7], "nvlink_exchange_s": flat[8],
            "preliminary": {"piles": len(walk["preliminary"]), "graphs_s": flat[2], "candidates_s": flat[3],
                            "metacells": int(walk.get("preliminary_metacells", walk["metacells"]))},
            "groups_s": flat[4],
            "final": {"piles": len(final), "pile_cells": [int(min(sizes)), int(max(sizes))], "graphs_s": flat[5],
                      "candidates_s": flat[6], "metacells": int(walk["metacells"])},
            "knn_graphs_s": flat[2] + flat[5],
            "graph_edges": int(sum(pile["edges"] for pile in walk["preliminary"]) + sum(pile["edges"] for pile in walk.get("final", []))),
            "h2d_bytes_this_rank": int(upload["h2d_bytes"]), "nvlink_bytes_this_rank": int(upload["exchanged_bytes"]),
            "target_pile_size": walk["target_pile_size"], "generate_s_untimed": generate_s,
            "pile_steps": args.atlas_pile_steps,
            "outlier_cells": int((walk["metacell_of_cells"] < 0).sum()),
            "data": "synthetic, distinct cells (tests/synth.py's model, generated per rank on the GPU, moved to page-locked host "
                    "memory before the timed region); max over ranks, barrier on both sides; wall_s = atlas upload + exchange + "
                    "cell totals + preliminary phase + groups + final phase; candidates = choose_seeds + first "
                    "optimize_partitions of every pile, side by side, one SM each"}


def time_oversized(args, rank, world, device, barrier, genes=2000, top=150):
    """The one exchange step of the path: the similarity of a pile too large for one GPU's N x N matrices,
    sharded by row blocks over the ranks (every rank holds the whole [cells, genes] operand and computes its rows on the
    tensor cores), the `top` largest values of every row kept, and only those gathered: one NCCL all-gather of
    cells x top indices and values.  Host array in, CSR out on every rank; the best of three, max over ranks."""
    import torch
    import torch.distributed as dist

    from metacells_b200 import tools as tl

    cells = args.oversized_cells
    rng = np.random.default_rng(SEED)
    types = rng.normal(size=(64, genes)).astype("float32")
    dense = np.log2(rng.poisson(np.exp(1.0 + 0.6 * types[rng.integers(0, 64, size=cells)])).astype("float32") + np.float32(1 / 7))
    walls = []
    for _ in range(3):
        barrier()
        start = time.perf_counter()
        matrix = tl.compute_obs_obs_similarity_top(dense, top)
        torch.cuda.synchronize(device)
        elapsed = torch.tensor([time.perf_counter() - start], dtype=torch.float64, device=device)
        dist.all_reduce(elapsed, op=dist.ReduceOp.MAX)
        walls.append(float(elapsed.item()))
    return {"metric": "oversized_pile_similarity_top_wall_time", "cells": int(cells), "genes": int(genes), "top": int(top),
            "n_gpus": world, "wall_s": min(walls), "wall_s_of_runs": walls, "nnz": int(matrix.nnz),
            "all_gather_bytes": int(cells) * top * 8, "rows_per_rank": int(-(-cells // world)),
            "note": "tools.compute_obs_obs_similarity_top: H2D of the operand, row-block similarity (tcgen05) + per-row top-k per rank, "
                    "one all-gather of indices and one of values over NVLink, CSR assembled on the host of every rank"}


def time_candidates(args, umis):
    """What follows the k-NN graph in a pile (tools/candidates.py:128-196, 351-368): choose_seeds and the first
    optimize_partitions, on the pile's graph as it lies on the device.  Both are one persistent CTA, so the piles of a
    worker run side by side, one SM and one host thread each; every pile here is the bench pile's graph."""
    from concurrent.futures import ThreadPoolExecutor

    from metacells_b200 import pipeline

    cells = umis.shape[0]
    graph, cell_umis = pipeline.pile_knn_graph_on_device(umis, features=args.features, random_seed=SEED)
    bounds = pipeline.metacell_bounds()
    wanted = pipeline.seeds_count_for(float(sum(cell_umis)), cells, bounds)

    def candidates(_=None):
        return graph.seed_partition(cell_umis, bounds, wanted, random_seed=SEED)

    try:
        candidates()  # warm-up: pools, attributes
        start = time.perf_counter()
        communities, score, seeds = candidates()
        alone_s = time.perf_counter() - start
        with ThreadPoolExecutor(max_workers=args.candidate_piles) as pool:
            start = time.perf_counter()
            results = list(pool.map(candidates, range(args.candidate_piles)))
            together_s = time.perf_counter() - start
        # the whole loop (tools/candidates.py:198-289: re-seeding, splits, minimal cuts) and the deviants pass that follows
        # it in the pile (pipeline/direct.py:311-322), once each
        from metacells_b200.candidates import compute_candidate_metacells
        from metacells_b200.deviants import find_deviant_cells

        rounds = []
        start = time.perf_counter()
        candidate = compute_candidate_metacells(graph, cell_umis, random_seed=SEED, trace=rounds)
        loop_s = time.perf_counter() - start
        with pipeline.Pile(umis) as pile:
            gaps = []
            start = time.perf_counter()
            deviants = find_deviant_cells(pile, candidate, total_umis_per_cell=cell_umis, trace=gaps)
            deviants_s = time.perf_counter() - start
    finally:
        graph.close()
    identical = all(np.array_equal(communities, other[0]) and score == other[1] for other in results)
    whole = {"candidates_loop_s": loop_s, "optimisations": sum(step[0] == "optimize" for step in rounds),
             "candidates": int(candidate.max()) + 1, "deviants_s": deviants_s, "deviants_rounds": len(gaps),
             "deviant_cells": int(deviants.sum()),
             "cpu_measured": "profiles/r02_candidates_deviants_16000.json: the oracle's loop 45.0 s (one thread, as the "
                             "reference's natives are), the reference's compute_cell_gaps loop 3.1 s (16 threads); == both"}
    return {"whole_loop_and_deviants": whole, "metric": "seeds_and_first_partition_optimisation_wall_time", "cells_per_pile": int(cells), "seeds": int(seeds),
            "one_pile_s": alone_s, "piles_side_by_side": int(args.candidate_piles), "side_by_side_s": together_s,
            "s_per_pile_side_by_side": together_s / args.candidate_piles, "identical_results": bool(identical),
            "note": "mcb_graph_seed_partition on the pile's k-NN graph as it lies on the device (incoming layout, choose_seeds, "
                    "optimize_partitions; cell UMIs up, partition down); one CTA = one SM per pile, one host thread and stream "
                    "per pile"}


def measure_tf32_peak(device):
    """Dense TF32 throughput of this GPU: the denominator of the contraction's roofline line, nothing else."""
    import torch

    previous = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    try:
        size = 8192
        left = torch.randn(size, size, device=device)
        right = torch.randn(size, size, device=device)
        out = torch.empty(size, size, device=device)
        for _ in range(3):
            torch.matmul(left, right, out=out)
        torch.cuda.synchronize(device)
        flops = 2.0 * size ** 3
        best = float("inf")
        for _ in range(10):
            begin, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            begin.record()
            torch.matmul(left, right, out=out)
            end.record()
            end.synchronize()
            best = min(best, begin.elapsed_time(end))
        begin, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        count = max(10, int(2000.0 / best))  # about two seconds back to back
        begin.record()
        for _ in range(count):
            torch.matmul(left, right, out=out)
        end.record()
        end.synchronize()
        sustained = begin.elapsed_time(end) / count
        return {"burst_tflops": flops / (best * 1e-3) / 1e12, "sustained_tflops": flops / (sustained * 1e-3) / 1e12,
                "how": "torch.matmul float32 with allow_tf32, 8192^3: best of 10 (burst), back to back for ~2 s (sustained)"}
    finally:
        torch.backends.cuda.matmul.allow_tf32 = previous


def ncu_traffic(kernel, cells, features):
    """DRAM bytes per launch of a kernel from the committed `ncu --set full` captures (profiles/r02_ncu_full_*.txt and the
    table of profiles/README.md, round 2; correlate at 2 000 genes: profiles/r01_ncu_full_correlate_umma.txt); null for a
    workload that was not captured."""
    if cells != 16000:
        return None
    table = {"rank_matrix": 1.025e9 + 0.969e9, "downsample": 0.690e9 + 0.206e9, "balance": 1.151e9 + 0.179e9 + 0.565e9 + 0.117e9}
    if features == 2000:
        table["correlate_umma"] = 3.302457e9 + 1.005782e9
    elif features == 890:
        table["correlate_umma"] = 0.998e9 + 0.994e9
    return table.get(kernel)


_RESULT_FD = None


def emit(line):
    """The one JSON line of the contract, on the process's real stdout."""
    text = (json.dumps(line) + "\n").encode()
    if _RESULT_FD is None:
        sys.stdout.write(text.decode())
        sys.stdout.flush()
    else:
        os.write(_RESULT_FD, text)


def main():
    global _RESULT_FD
    # Libraries print on stdout (NCCL's version banner, for one): keep fd 1 for the JSON line only.
    sys.stdout.flush()
    _RESULT_FD = os.dup(1)
    os.dup2(2, 1)
    args = parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_gpu_arm(args)


if __name__ == "__main__":
    main()
