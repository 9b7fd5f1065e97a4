"""Developer diagnostic (GPU box): end-to-end step time (pinned host pile in, graph out) against the number of
upload chunks (MCB_PILE_CHUNKS), several rounds interleaved inside one process."""
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
pinned = {n: torch.from_numpy(a).pin_memory() for n, a in (("data", umis.data), ("indices", umis.indices), ("indptr", umis.indptr))}
p = bench.p
def step():
    pile = ctypes.c_void_p()
    lib.call("mcb_pile_create_host", p(pinned["data"].data_ptr()), p(pinned["indices"].data_ptr()), p(pinned["indptr"].data_ptr()), ctypes.c_int64(cells), ctypes.c_int64(genes), ctypes.byref(pile))
    lib.call("mcb_pile_set_max_cell_total", pile, ctypes.c_uint64(bench.make_pile.max_cell_total))
    lib.call("mcb_pile_downsample", pile, ctypes.c_int32(samples), ctypes.c_uint64(1))
    lib.call("mcb_pile_top_genes", pile, ctypes.c_int64(args.features), None)
    graph = ctypes.c_void_p(); outside = ctypes.c_int(0)
    lib.call("mcb_pile_knn_graph", pile, None, ctypes.c_int64(args.features), p(table.ctypes.data), ctypes.c_int64(len(table)), ctypes.c_float(np.float32(1/7)), ctypes.c_int(1), ctypes.c_int64(k), ctypes.c_double(tl.KNN_BALANCED_RANKS_FACTOR), ctypes.c_double(3.0), ctypes.c_double(1.0), ctypes.c_int64(2), ctypes.byref(outside), ctypes.byref(graph))
    lib.lib.mcb_graph_free(graph); lib.lib.mcb_pile_free(pile)
for _ in range(3): step()
for round_ in range(3):
    for chunks in (1, 2, 4, 8, 16):
        os.environ["MCB_PILE_CHUNKS"] = str(chunks)
        step(); torch.cuda.synchronize(); t0 = time.perf_counter()
        for _ in range(8): step()
        torch.cuda.synchronize()
        print("round", round_, "chunks", chunks, "%.2f ms" % ((time.perf_counter() - t0) / 8 * 1e3), flush=True)
