"""Developer diagnostic (GPU box): the steps after the k-NN graph of one synthetic pile, GPU against the CPU --
the whole compute_candidate_metacells loop (metacells_b200.candidates) and find_deviant_cells (metacells_b200.deviants).
  python scripts/time_candidates.py [--cells 16000] [--genes 27000] [--features select] [--cpu]
--cpu also runs the oracle's loop (the C restatement of the two natives, one thread) and the reference's
compute_cell_gaps (oracle/_ref, all host threads) on the same inputs and checks `==`."""
import argparse, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from synth import synth_umis
from metacells_b200 import _lib, pipeline
from metacells_b200.candidates import compute_candidate_metacells
from metacells_b200.deviants import find_deviant_cells

parser = argparse.ArgumentParser()
parser.add_argument("--cells", type=int, default=16000)
parser.add_argument("--genes", type=int, default=27000)
parser.add_argument("--features", default="select")
parser.add_argument("--cpu", action="store_true")
parser.add_argument("--out", default="")
args = parser.parse_args()
SEED = 123456
features = args.features if args.features == "select" else int(args.features)
report = {"cells": args.cells, "genes": args.genes, "features": args.features}
umis = synth_umis(args.cells, args.genes, seed=SEED)
_lib.call("mcb_profile_enable", ("int", 1))


def kernel_times():
    import ctypes
    library = _lib.lib()
    buffer = ctypes.create_string_buffer(8192)
    library.mcb_profile_names(buffer, ctypes.c_longlong(8192))
    times = {}
    for name in buffer.value.decode().split("\n"):
        if not name:
            continue
        total, count = ctypes.c_double(0), ctypes.c_longlong(0)
        library.mcb_profile_read(name.encode(), ctypes.byref(total), ctypes.byref(count))
        if count.value:
            times[name] = (round(total.value, 3), count.value)
    library.mcb_profile_reset()
    return times


start = time.perf_counter()
graph, cell_umis = pipeline.pile_knn_graph_on_device(umis, features=features, random_seed=SEED)
report["knn_graph_s"] = round(time.perf_counter() - start, 3)
kernel_times()
trace = []
start = time.perf_counter()
candidate = compute_candidate_metacells(graph, cell_umis, random_seed=SEED, trace=trace)
report["candidates_gpu_s"] = round(time.perf_counter() - start, 3)
report["candidates_steps"] = [step[0] for step in trace]
report["candidates_count"] = int(candidate.max()) + 1
report["candidates_kernels_ms"] = kernel_times()
print(json.dumps(report), flush=True)

with pipeline.Pile(umis) as pile:
    gaps = []
    start = time.perf_counter()
    deviants = find_deviant_cells(pile, candidate, total_umis_per_cell=cell_umis, trace=gaps)
    report["deviants_gpu_s"] = round(time.perf_counter() - start, 3)
report["deviants_rounds"] = len(gaps)
report["deviants_found"] = int(deviants.sum())
report["deviants_kernels_ms"] = kernel_times()
print(json.dumps(report), flush=True)

if args.cpu:
    from oracle import oracle

    outgoing, _ = graph.to_scipy()
    start = time.perf_counter()
    expected = oracle.compute_candidate_metacells(outgoing, cell_umis, random_seed=SEED)
    report["candidates_cpu_oracle_s"] = round(time.perf_counter() - start, 3)
    report["candidates_equal"] = bool(np.array_equal(candidate, expected))
    print(json.dumps(report), flush=True)
    ref = oracle.load_ref()
    if ref is not None and hasattr(ref, "compute_cell_gaps"):
        theirs = []
        start = time.perf_counter()
        expected = oracle.gaps_deviant_cells(umis, candidate, ref, trace=theirs)
        report["deviants_cpu_reference_s"] = round(time.perf_counter() - start, 3)
        report["deviants_equal"] = bool(np.array_equal(deviants, expected) and len(gaps) == len(theirs)
                                        and all(np.array_equal(a, b) for a, b in zip(gaps, theirs)))
graph.close()
print(json.dumps(report), flush=True)
if args.out:
    with open(args.out, "w") as handle:
        json.dump(report, handle, indent=1)
