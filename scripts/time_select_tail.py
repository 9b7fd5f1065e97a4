"""Developer diagnostic (GPU box): where the host side of the feature-gene selection spends its time on bench.py's pile.
   python scripts/time_select_tail.py"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import bench
from metacells_b200 import pipeline, select

umis, samples, k = bench.make_pile(16000, 27000, bench.SEED)
laps = {}


def lap(name, start):
    laps.setdefault(name, []).append((time.perf_counter() - start) * 1e3)


with pipeline.Pile(umis, max_cell_total=bench.make_pile.max_cell_total) as pile:
    pile.wait_upload()
    pile.want_gene_statistics(select.SELECT_MIN_GENE_TOP3)
    pile.downsample(samples, bench.SEED)
    for attempt in range(30):
        start = time.perf_counter(); statistics = pile.gene_statistics(); lap("gene_statistics (D2H)", start)
        start = time.perf_counter(); sums, squares, tops = select.gene_sums(pile); lap("gene_sums (incl. statistics + columns)", start)
        start = time.perf_counter(); order = np.argsort(sums / 16000); lap("argsort", start)
        values = np.log2(np.maximum(sums, 1)).astype("float32")
        start = time.perf_counter(); select._device_ordered_medians(values, order, 101); lap("ordered medians call", start)
        start = time.perf_counter(); relative = select.relative_variance_of_genes(sums, squares, 16000); lap("relative_variance_of_genes", start)
        start = time.perf_counter(); mask = select.select_feature_genes(pile); lap("select_feature_genes (all)", start)
for name, values in laps.items():
    print("%-44s %.3f ms (median of %d)" % (name, float(np.median(values[5:])), len(values) - 5))
print(int(mask.sum()), "genes")
