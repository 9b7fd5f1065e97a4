"""Feature-gene selection on bench.py's pile (GPU box): the cost of the statistics in the downsample kernel and of the
host tail, and the genes selected.   python scripts/time_select.py [--cells 16000]"""
import argparse, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import bench
from metacells_b200 import pipeline, select

parser = argparse.ArgumentParser()
parser.add_argument("--cells", type=int, default=16000)
parser.add_argument("--genes", type=int, default=27000)
args = parser.parse_args()
umis, samples, k = bench.make_pile(args.cells, args.genes, bench.SEED)
for attempt in range(4):
    for with_statistics in (False, True):
        with pipeline.Pile(umis, max_cell_total=bench.make_pile.max_cell_total) as pile:
            pile.wait_upload()
            if with_statistics:
                pile.want_gene_statistics(select.SELECT_MIN_GENE_TOP3)
            start = time.perf_counter()
            pile.downsample(samples, bench.SEED)
            pile.gene_totals()  # synchronises
            downsampled = time.perf_counter() - start
            if with_statistics:
                start = time.perf_counter()
                mask = select.select_feature_genes(pile)
                selected = time.perf_counter() - start
                start = time.perf_counter()
                sums = select.gene_sums(pile)
                summed = time.perf_counter() - start
                print(f"downsample + statistics {downsampled * 1e3:.2f} ms, selection {selected * 1e3:.2f} ms (sums {summed * 1e3:.2f} ms), "
                      f"{int(mask.sum())} genes selected")
            else:
                print(f"downsample alone {downsampled * 1e3:.2f} ms")
