"""rank_matrix on the similarity matrix of a synthetic pile (GPU box): the kernel's time on the distribution it
meets in the path, and the bucket statistics of a few rows (a numpy emulation of the kernel's binning).
  python scripts/rank_on_pile.py [--cells 16000] [--repeat 5]        (MCB_RANK_MERGE=1 for the merge sort)"""
import argparse, ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
from metacells_b200 import _lib, pipeline, tools as tl
from synth import synth_umis

parser = argparse.ArgumentParser()
parser.add_argument("--cells", type=int, default=16000)
parser.add_argument("--genes", type=int, default=27000)
parser.add_argument("--features", type=int, default=2000)
parser.add_argument("--repeat", type=int, default=5)
args = parser.parse_args()
library = _lib.lib()
p = lambda tensor: ctypes.c_void_p(tensor.data_ptr())
i64 = ctypes.c_int64
umis = synth_umis(args.cells, args.genes, seed=123456)
totals = np.asarray(umis.sum(axis=1)).reshape(-1).astype("float32")
samples = int(round(min(max(tl.DOWNSAMPLE_MIN_SAMPLES, np.quantile(totals, tl.DOWNSAMPLE_MIN_CELL_QUANTILE)), np.quantile(totals, tl.DOWNSAMPLE_MAX_CELL_QUANTILE))))
table = pipeline.log_fraction_table(samples)
similarity = torch.empty((args.cells, args.cells), dtype=torch.float32, device="cuda")
with pipeline.Pile(umis) as pile:
    pile.downsample(samples, 123456)
    pile.select_top_genes(args.features)
    outside = ctypes.c_int(0)
    _lib.check(library.mcb_pile_similarity(pile._handle, None, i64(args.features), ctypes.c_void_p(table.ctypes.data), i64(len(table)),
                                           ctypes.c_float(np.float32(1 / 7)), ctypes.c_int(1), p(similarity), ctypes.byref(outside)))
torch.cuda.synchronize()
similarity.fill_diagonal_(float(similarity.min()) - 1)
work = torch.empty_like(similarity)
stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)

def read(name):
    total, calls = ctypes.c_double(0), ctypes.c_longlong(0)
    library.mcb_profile_read(name.encode(), ctypes.byref(total), ctypes.byref(calls))
    return total.value / max(calls.value, 1)

for attempt in range(args.repeat + 1):
    if attempt == 1:
        library.mcb_profile_reset(); library.mcb_profile_enable(ctypes.c_int(1))
    work.copy_(similarity)
    _lib.check(library.mcb_dev_rank_matrix_f32(p(work), i64(args.cells), i64(args.cells), i64(args.cells), ctypes.c_int(0), stream))
torch.cuda.synchronize()
print("rank_matrix ms", read("rank_matrix"), "ties ms", read("rank_matrix_ties"))
# parity with torch: descending ranks, ties by position
rows = [0, 17, args.cells // 2, args.cells - 1]
for row in rows:
    values = similarity[row]
    order = torch.sort(values, descending=True, stable=True).indices
    expected = torch.empty(args.cells, device="cuda"); expected[order] = torch.arange(1, args.cells + 1, device="cuda", dtype=torch.float32)
    assert torch.equal(expected, work[row]), row
# bucket statistics
N = args.cells; BINS = 16384 if N > 8192 else (8192 if N > 4096 else 4096)
for row in rows:
    v = similarity[row].cpu().numpy()
    srt = np.sort(v); largest = srt[-1]; low = srt[1]
    scale = np.float32(1024) / np.float32(largest - low)
    t = np.clip((np.float32(largest) - v) * scale, 0, np.float32(1023.9999)).astype("float32")
    c = t.astype(int); cnt = np.bincount(c, minlength=1024); base = np.cumsum(cnt) - cnt
    est = (base[c] + (t - c) * cnt[c]).astype("float32")
    fine = np.minimum((est * np.float32(BINS / N)).astype(int), BINS - 1)
    fc = np.bincount(fine, minlength=BINS)
    print(f"row {row}: unique {len(np.unique(v))}/{N}, range [{low:.4f}, {largest:.4f}], quartiles {np.percentile(v, [25, 50, 75]).round(4)}, "
          f"coarse max {cnt.max()} nonempty {(cnt > 0).sum()}, fine max {fc.max()}, size-biased mean {(fc * fc).sum() / N:.2f}, sizes {np.bincount(fc)[:10]}")
