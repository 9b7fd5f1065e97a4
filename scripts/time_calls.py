"""Developer diagnostic (GPU box): wall time of every C-ABI call of one resident step."""
import os, sys, time, ctypes
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import bench
from metacells_b200 import pipeline, tools as tl

args = bench.parse_args()
lib = bench.Library()
umis, samples, k = bench.make_pile(args.cells, args.genes, bench.SEED)
cells, genes = umis.shape
table = pipeline.log_fraction_table(samples)
dev = {n: torch.from_numpy(a).cuda() for n, a in (("data", umis.data), ("indices", umis.indices), ("indptr", umis.indptr))}
totals = np.empty(genes, dtype="float32")
p = bench.p
def sync(): torch.cuda.synchronize()
for it in range(4):
    t = {}
    def lap(name, t0): sync(); t[name] = (time.perf_counter() - t0) * 1e3
    sync(); t_all = time.perf_counter()
    t0 = time.perf_counter(); pile = ctypes.c_void_p()
    lib.call("mcb_pile_create_dev", p(dev["data"].data_ptr()), p(dev["indices"].data_ptr()), p(dev["indptr"].data_ptr()), ctypes.c_int64(cells), ctypes.c_int64(genes), p(0), ctypes.byref(pile)); lap("create", t0)
    t0 = time.perf_counter(); lib.call("mcb_pile_downsample", pile, ctypes.c_int32(samples), ctypes.c_uint64(1)); lap("downsample", t0)
    t0 = time.perf_counter(); lib.call("mcb_pile_gene_totals", pile, p(totals.ctypes.data)); lap("totals", t0)
    t0 = time.perf_counter(); fg = np.sort(np.argsort(-totals, kind="stable")[:args.features]).astype("int32"); lap("argsort", t0)
    t0 = time.perf_counter(); sim = torch.empty((cells, cells), dtype=torch.float32, device="cuda"); outside = ctypes.c_int(0)
    lib.call("mcb_pile_similarity", pile, p(fg.ctypes.data), ctypes.c_int64(len(fg)), p(table.ctypes.data), ctypes.c_int64(len(table)), ctypes.c_float(np.float32(1/7)), ctypes.c_int(1), p(sim.data_ptr()), ctypes.byref(outside)); lap("similarity", t0)
    t0 = time.perf_counter(); graph = ctypes.c_void_p()
    lib.call("mcb_dev_knn_graph_f32", p(sim.data_ptr()), ctypes.c_int64(cells), ctypes.c_int64(k), ctypes.c_double(tl.KNN_BALANCED_RANKS_FACTOR), ctypes.c_double(3.0), ctypes.c_double(1.0), ctypes.c_int64(2), p(0), ctypes.byref(graph)); lap("knn_graph", t0)
    t0 = time.perf_counter(); lib.lib.mcb_graph_free(graph); lib.lib.mcb_pile_free(pile); lap("free", t0)
    sync(); total = (time.perf_counter() - t_all) * 1e3
    print(it, "total %.1f" % total, {n: round(v, 2) for n, v in t.items()}, flush=True)
