"""Developer diagnostic (GPU box): the divide-and-conquer walk over a small synthetic atlas; when a pile's seeds + first
optimisation refuse the graph, says what the graph looks like.   python scripts/diag_small_atlas.py [cells]"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import bench
from metacells_b200 import divide_and_conquer as dc, tools
from metacells_b200._lib import McbError

cells = int(sys.argv[1]) if len(sys.argv) > 1 else 60000
device = torch.device("cuda", 0)
original = tools.KnnGraph.seed_partition


def verbose(self, cell_umis, bounds, wanted, **arguments):
    try:
        return original(self, cell_umis, bounds, wanted, **arguments)
    except McbError as error:
        outgoing, _ = self.to_scipy()
        sums = np.asarray(outgoing.sum(axis=1)).reshape(-1).astype("float64")
        counts = np.diff(outgoing.indptr)
        print(f"[diag] {error}\n[diag] nodes {outgoing.shape[0]} k {self.k} nnz {outgoing.nnz} wanted seeds {wanted} "
              f"row entries min {counts.min()} max {counts.max()} empty rows {(counts == 0).sum()} "
              f"row sums min {sums.min():.9f} max {sums.max():.9f} off by > 1e-6: {(np.abs(sums - 1) > 1e-6).sum()} "
              f"nan weights {np.isnan(outgoing.data).sum()} masses min {np.min(cell_umis)} max {np.max(cell_umis)}", flush=True)
        worst = int(np.argmax(np.abs(sums - 1)))
        print(f"[diag] worst row {worst}: entries {counts[worst]} weights {outgoing.data[outgoing.indptr[worst]:outgoing.indptr[worst + 1]][:12]}", flush=True)
        raise


tools.KnnGraph.seed_partition = verbose
data, indices, indptr = bench.synth_atlas_share(cells, 27000, bench.SEED, 0, cells, device)
atlas = dc.Atlas.from_host_shares(data, indices, indptr, 27000, rank=0, world_size=1, device=device)
started = time.perf_counter()
try:
    walk = dc.divide_and_conquer_graphs(atlas, features="select", random_seed=bench.SEED, workers=2)
    print("[diag] walk ok", {name: len(walk[name]) for name in ("preliminary", "final") if name in walk}, round(time.perf_counter() - started, 2))
except McbError as error:
    print("[diag] walk failed after", round(time.perf_counter() - started, 2), "s:", error)
finally:
    atlas.close()
