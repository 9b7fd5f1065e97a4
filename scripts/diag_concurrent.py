"""Developer diagnostic (GPU box): are the one-CTA optimisations of several host threads running side by side?
Prints every thread's start / end of its optimize_partitions call and the kernels' own durations (CUDA events)."""
import argparse, ctypes, os, sys, threading, time
if "--connections" in sys.argv:
    os.environ["CUDA_DEVICE_MAX_CONNECTIONS"] = sys.argv[sys.argv.index("--connections") + 1]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, scipy.sparse as sp
from synth import synth_umis
from metacells_b200 import pipeline, extensions as xt, tools as tl, _lib

parser = argparse.ArgumentParser()
parser.add_argument("--cells", type=int, default=4000)
parser.add_argument("--genes", type=int, default=8000)
parser.add_argument("--threads", type=int, default=4)
parser.add_argument("--profile", type=int, default=1)
parser.add_argument("--reserve-mb", type=int, default=0)
parser.add_argument("--device-arrays", type=int, default=0, help="1: the graph stays on the device (mcb_dev_optimize_partitions, a torch stream per thread)")
parser.add_argument("--connections", type=int, default=0)
args = parser.parse_args()
SEED = 123456
umis = synth_umis(args.cells, args.genes, seed=SEED)
cell_umis = np.asarray(umis.sum(axis=1)).reshape(-1).astype("float32")
outgoing, details = pipeline.compute_pile_knn_graph(umis, features=1000, random_seed=SEED)
outgoing, incoming = tl._both_layouts(outgoing)
graph = (outgoing.data, outgoing.indices, outgoing.indptr, incoming.data, incoming.indices, incoming.indptr)
bounds = pipeline.metacell_bounds()
wanted = pipeline.seeds_count_for(float(sum(cell_umis)), args.cells, bounds)
seeds = np.full(args.cells, -1, dtype="int32")
count = xt.choose_seeds(*graph, SEED, wanted, 0.85, 0.95, seeds)
mask = np.ones(max(count, 2), dtype="int8")
_lib.call("mcb_profile_enable", ("int", args.profile))
if args.reserve_mb:
    _lib.call("mcb_reserve_scratch", ("i64", args.reserve_mb << 20))
origin = time.perf_counter()
if args.device_arrays:
    import torch
    device_graph = [torch.from_numpy(array).cuda() for array in graph]
    device_umis, device_mask = torch.from_numpy(cell_umis).cuda(), torch.from_numpy(mask).cuda()
    torch.cuda.synchronize()
def work(index, log):
    mine = seeds.copy()
    start = time.perf_counter() - origin
    if args.device_arrays:
        stream = torch.cuda.Stream()
        partition = torch.from_numpy(mine).cuda()
        torch.cuda.synchronize()
        start = time.perf_counter() - origin
        score = ctypes.c_double(0)
        pointer = lambda tensor: ctypes.c_void_p(tensor.data_ptr())
        function = _lib.lib().mcb_dev_optimize_partitions
        function.restype = ctypes.c_int
        _lib.check(function(*[pointer(tensor) for tensor in device_graph], ctypes.c_int64(args.cells), ctypes.c_uint32(SEED),
                            pointer(device_umis), *[ctypes.c_double(bound) for bound in bounds], ctypes.c_double(0.02),
                            ctypes.c_double(0.25), pointer(partition), pointer(device_mask), ctypes.c_int64(len(mask)),
                            ctypes.c_double(0.98), None, ctypes.byref(score), ctypes.c_void_p(stream.cuda_stream)))
        score = score.value
    else:
        score = xt.optimize_partitions(*graph, SEED, cell_umis, *bounds, 0.02, 0.25, mine, mask, 0.98)
    log[index] = (start, time.perf_counter() - origin, score)
for threads in (1, args.threads):
    origin = time.perf_counter()
    log = [None] * threads
    pool = [threading.Thread(target=work, args=(index, log)) for index in range(threads)]
    for thread in pool: thread.start()
    for thread in pool: thread.join()
    print("%d threads: wall %.3f s" % (threads, time.perf_counter() - origin))
    for index, (start, end, score) in enumerate(log):
        print("   thread %d: %.3f .. %.3f s  score %.6f" % (index, start, end, score))
if args.profile:
    library = _lib.lib()
    total, calls = ctypes.c_double(0), ctypes.c_longlong(0)
    library.mcb_profile_read(b"optimize_partitions", ctypes.byref(total), ctypes.byref(calls))
    print("optimize_partitions kernel: %.1f ms over %d launches (its own duration, CUDA events)" % (total.value, calls.value))
