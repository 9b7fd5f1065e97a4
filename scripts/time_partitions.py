"""Developer diagnostic (GPU box): optimize_partitions on the k-NN graph of one synthetic pile, GPU against
the reference's CPU implementation (oracle/_ref when it travelled, else the C oracle), same inputs and seed.
  python scripts/time_partitions.py [--cells 16000] [--genes 27000] [--features 2000] [--no-cpu]"""
import argparse, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, scipy.sparse as sp
from synth import synth_umis, metacell_bounds, seeds_count_for
from oracle import oracle
from metacells_b200 import pipeline, extensions as xt, _lib

parser = argparse.ArgumentParser()
parser.add_argument("--cells", type=int, default=16000)
parser.add_argument("--genes", type=int, default=27000)
parser.add_argument("--features", type=int, default=2000)
parser.add_argument("--no-cpu", action="store_true")
parser.add_argument("--repeat", type=int, default=2)
parser.add_argument("--k", type=int, default=0, help="k of the k-NN graph (0: from the pile, direct.py:251-267)")
args = parser.parse_args()
SEED = 123456
ref = oracle.load_ref()
umis = synth_umis(args.cells, args.genes, seed=SEED)
cell_umis = np.asarray(umis.sum(axis=1)).reshape(-1).astype("float32")
start = time.perf_counter()
outgoing, details = pipeline.compute_pile_knn_graph(umis, features=args.features, random_seed=SEED, k=args.k or None)
print("knn graph on the GPU: %.2f s, k=%d, nnz=%d" % (time.perf_counter() - start, details["k"], outgoing.nnz), flush=True)
incoming = sp.csc_matrix(outgoing); incoming.sort_indices()
graph = (outgoing.data, outgoing.indices, outgoing.indptr, incoming.data, incoming.indices, incoming.indptr)
bounds = metacell_bounds(48, 160000)
count = seeds_count_for(float(cell_umis.sum()), args.cells, bounds)
initial = np.full(args.cells, -1, dtype="int32")
if ref is not None:
    start = time.perf_counter()
    count = ref.choose_seeds(*graph, SEED, count, 0.8, 1.0, initial)
    print("reference choose_seeds: %.2f s, %d seeds" % (time.perf_counter() - start, count), flush=True)
else:
    initial = np.random.default_rng(SEED).permutation(np.arange(args.cells) % count).astype("int32")
mask = np.ones(max(count, 2), dtype="int8")
_lib.call("mcb_profile_enable", ("int", 1))
for it in range(args.repeat):
    mine = initial.copy()
    start = time.perf_counter()
    my_score = xt.optimize_partitions(*graph, SEED, cell_umis, *bounds, 0.02, 0.25, mine, mask, 0.98)
    elapsed = time.perf_counter() - start
    stats = xt.last_optimize_stats.copy()
    print("GPU optimize_partitions: %.3f s score %.17g passes %d visits %d moves %d skipped %d batches %d" % (elapsed, my_score, *stats[:5]), flush=True)
    print("   Mcycles: setup %.0f evaluate %.0f choose %.0f move %.0f | warp0 scan+neighbours %.0f candidates %.0f" % tuple(stats[8:14] / 1e6), flush=True)
    print("   warp0 Mcycles: partition scan %.0f, gathers + re-scoring %.0f, ordered sums %.0f" % (stats[6] / 1e6, stats[7] / 1e6, stats[14] / 1e6), flush=True)
if not args.no_cpu:
    theirs = initial.copy()
    start = time.perf_counter()
    if ref is not None:
        their_score = ref.optimize_partitions(*graph, SEED, cell_umis, *bounds, 0.02, 0.25, theirs, mask, 0.98)
    else:
        their_score = oracle.optimize_partitions(outgoing, incoming, SEED, cell_umis, *bounds, 0.02, 0.25, theirs, mask, 0.98)
    print("CPU %s optimize_partitions: %.3f s score %.17g equal partitions %s equal score %s" % (
        "reference" if ref is not None else "oracle", time.perf_counter() - start, their_score, np.array_equal(mine, theirs), my_score == their_score), flush=True)
import ctypes
library = _lib.lib()
buffer = ctypes.create_string_buffer(4096)
library.mcb_profile_names(buffer, ctypes.c_longlong(4096))
for name in buffer.value.decode().split("\n"):
    if not name:
        continue
    total, calls = ctypes.c_double(0), ctypes.c_longlong(0)
    library.mcb_profile_read(name.encode(), ctypes.byref(total), ctypes.byref(calls))
    if calls.value:
        print("  kernel %-24s %10.3f ms over %d launches" % (name, total.value, calls.value))
