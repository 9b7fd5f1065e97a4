"""Developer diagnostic (GPU box): choose_seeds on the k-NN graph of one synthetic pile, GPU against the reference's
CPU implementation (oracle/_ref when it travelled, else the C oracle), same inputs and seed.
  python scripts/time_seeds.py [--cells 16000] [--genes 27000] [--features 2000]"""
import argparse, ctypes, os, sys, time
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
parser.add_argument("--repeat", type=int, default=3)
args = parser.parse_args()
SEED = 123456
ref = oracle.load_ref()
umis = synth_umis(args.cells, args.genes, seed=SEED)
cell_umis = np.asarray(umis.sum(axis=1)).reshape(-1).astype("float32")
outgoing, details = pipeline.compute_pile_knn_graph(umis, features=args.features, random_seed=SEED)
incoming = sp.csc_matrix(outgoing); incoming.sort_indices()
graph = (outgoing.data, outgoing.indices, outgoing.indptr, incoming.data, incoming.indices, incoming.indptr)
bounds = metacell_bounds(48, 160000)
count = seeds_count_for(float(cell_umis.sum()), args.cells, bounds)
print("graph: %d nodes, %d edges, k=%d, %d seeds wanted" % (args.cells, outgoing.nnz, details["k"], count), flush=True)
_lib.call("mcb_profile_enable", ("int", 1))
for low, high in ((0.85, 0.95), (0.8, 1.0)):
    for it in range(args.repeat):
        mine = np.full(args.cells, -1, dtype="int32")
        start = time.perf_counter()
        my_count = xt.choose_seeds(*graph, SEED, count, low, high, mine)
        elapsed = time.perf_counter() - start
        print("GPU choose_seeds(%.2f, %.2f): %.1f ms host to host, %d seeds" % (low, high, elapsed * 1e3, my_count), flush=True)
    theirs = np.full(args.cells, -1, dtype="int32")
    start = time.perf_counter()
    if ref is not None:
        their_count = ref.choose_seeds(*graph, SEED, count, low, high, theirs)
    else:
        their_count = oracle.choose_seeds(outgoing, incoming, SEED, count, low, high, theirs)
    print("CPU %s choose_seeds: %.1f ms, equal seeds %s equal count %s" % (
        "reference" if ref is not None else "oracle", (time.perf_counter() - start) * 1e3, np.array_equal(mine, theirs),
        my_count == their_count), flush=True)
library = _lib.lib()
buffer = ctypes.create_string_buffer(4096)
library.mcb_profile_names(buffer, ctypes.c_longlong(4096))
for name in buffer.value.decode().split("\n"):
    if not name:
        continue
    total, calls = ctypes.c_double(0), ctypes.c_longlong(0)
    library.mcb_profile_read(name.encode(), ctypes.byref(total), ctypes.byref(calls))
    if calls.value and "seeds" in name:
        print("  kernel %-24s %10.3f ms over %d launches" % (name, total.value, calls.value))
