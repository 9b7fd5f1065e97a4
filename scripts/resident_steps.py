"""A few device-resident steps of bench.py's pile, nothing else: the command the `ncu --set full` captures of the path's
kernels run (profiles/README.md).   python scripts/resident_steps.py [--steps 3] [--cells 16000]"""
import argparse, ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import bench
from metacells_b200 import pipeline, select, tools as tl

parser = argparse.ArgumentParser()
parser.add_argument("--steps", type=int, default=3)
parser.add_argument("--cells", type=int, default=16000)
parser.add_argument("--genes", type=int, default=27000)
parser.add_argument("--features", default="select", type=lambda t: t if t == "select" else int(t))
parser.add_argument("--times", action="store_true", help="print the mean time of every kernel over the steps after the first two")
args = parser.parse_args()
library = bench.Library()
umis, samples, k = bench.make_pile(args.cells, args.genes, bench.SEED)
table = pipeline.log_fraction_table(samples)
device = torch.device("cuda", 0)
resident = {name: torch.from_numpy(array).to(device) for name, array in (("data", umis.data), ("indices", umis.indices), ("indptr", umis.indptr))}
stream = torch.cuda.current_stream(device)
p = bench.p
for step in range(args.steps):
    if args.times and step == 2:
        torch.cuda.synchronize()
        library.call("mcb_profile_enable", ctypes.c_int(1))
        library.call("mcb_profile_reset")
        started = torch.cuda.Event(enable_timing=True); started.record(stream)
    pile = ctypes.c_void_p()
    library.call("mcb_pile_create_dev", p(resident["data"].data_ptr()), p(resident["indices"].data_ptr()), p(resident["indptr"].data_ptr()),
                 ctypes.c_int64(args.cells), ctypes.c_int64(args.genes), p(stream.cuda_stream), ctypes.byref(pile))
    library.call("mcb_pile_set_max_cell_total", pile, ctypes.c_uint64(bench.make_pile.max_cell_total))
    if args.features == "select":
        library.call("mcb_pile_want_gene_statistics", pile, ctypes.c_float(select.SELECT_MIN_GENE_TOP3))
    library.call("mcb_pile_downsample", pile, ctypes.c_int32(samples), ctypes.c_uint64(bench.SEED))
    if args.features == "select":
        view = pipeline.Pile._adopt(pile, (args.cells, args.genes))
        view.gene_top_threshold = float(select.SELECT_MIN_GENE_TOP3)
        chosen = np.flatnonzero(select.select_feature_genes(view)).astype("int32")
        view._handle = None
        genes_pointer, genes_count = p(chosen.ctypes.data), len(chosen)
    else:
        library.call("mcb_pile_top_genes", pile, ctypes.c_int64(args.features), None)
        genes_pointer, genes_count = None, args.features
    graph = ctypes.c_void_p()
    library.call("mcb_pile_knn_graph", pile, genes_pointer, ctypes.c_int64(genes_count), p(table.ctypes.data), ctypes.c_int64(len(table)),
                 ctypes.c_float(np.float32(1 / 7)), ctypes.c_int(1), ctypes.c_int64(k), ctypes.c_double(tl.KNN_BALANCED_RANKS_FACTOR),
                 ctypes.c_double(tl.KNN_INCOMING_DEGREE_FACTOR), ctypes.c_double(tl.KNN_OUTGOING_DEGREE_FACTOR),
                 ctypes.c_int64(tl.KNN_MIN_OUTGOING_DEGREE), None, ctypes.byref(graph))
    print("step", step, "graph nnz", library.lib.mcb_graph_nnz(graph), "k", library.lib.mcb_graph_k(graph))
    library.lib.mcb_graph_free(graph)
    library.lib.mcb_pile_free(pile)
torch.cuda.synchronize()
if args.times and args.steps > 2:
    finished = torch.cuda.Event(enable_timing=True); finished.record(stream); torch.cuda.synchronize()
    print("ms per step (events enabled): %.3f" % (started.elapsed_time(finished) / (args.steps - 2)))
    buffer = ctypes.create_string_buffer(8192)
    library.lib.mcb_profile_names(buffer, ctypes.c_longlong(8192))
    for name in buffer.value.decode().split("\n"):
        if name:
            total, count = ctypes.c_double(0), ctypes.c_longlong(0)
            library.lib.mcb_profile_read(name.encode(), ctypes.byref(total), ctypes.byref(count))
            if count.value:
                print("  %-28s %8.3f ms per step over %d launches" % (name, total.value / (args.steps - 2), count.value))
