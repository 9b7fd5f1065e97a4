"""Developer diagnostic (GPU box): choose_seeds + the first optimize_partitions of P piles side by side (one CTA, one
host thread and stream per pile) against one pile at a time.  Every pile is the same synthetic pile's k-NN graph (the
work depends on sizes, not values), with its own random seed so the optimisations differ.
  python scripts/time_piles_candidates.py [--cells 16000] [--genes 27000] [--piles 1,8,32,63]"""
import argparse, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, scipy.sparse as sp
from synth import synth_umis
from metacells_b200 import pipeline

parser = argparse.ArgumentParser()
parser.add_argument("--cells", type=int, default=16000)
parser.add_argument("--genes", type=int, default=27000)
parser.add_argument("--features", type=int, default=2000)
parser.add_argument("--piles", default="1,8,32,63")
args = parser.parse_args()
SEED = 123456
umis = synth_umis(args.cells, args.genes, seed=SEED)
cell_umis = np.asarray(umis.sum(axis=1)).reshape(-1).astype("float32")
outgoing, details = pipeline.compute_pile_knn_graph(umis, features=args.features, random_seed=SEED)
print("graph: %d nodes, %d edges, k=%d" % (args.cells, outgoing.nnz, details["k"]), flush=True)
first = None
for count in [int(text) for text in args.piles.split(",")]:
    piles = [(outgoing, cell_umis)] * count
    start = time.perf_counter()
    results = pipeline.compute_piles_seed_partitions(piles, random_seed=SEED)
    elapsed = time.perf_counter() - start
    if first is None:
        first = results[0]
    same = all(np.array_equal(result[0], first[0]) and result[1] == first[1] for result in results)
    print("%3d piles side by side: %7.2f s wall, %6.2f s per pile, %d seeds, score %.6f, all equal to the first run: %s" % (
        count, elapsed, elapsed / count, results[0][2], results[0][1], same), flush=True)
