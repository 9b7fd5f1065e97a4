"""Kernel sweep of BASELINE.json config 2 (GPU box): correlate + rank_matrix + collect_top + collect_pruned over
cells x feature genes, device-resident, CUDA-event times through the library's profiler (mcb_profile_*).
  python scripts/sweep_kernels.py [--cells 2000,4000,...] [--genes 1000,2000,4000] [--repeat 3]
Prints one markdown table row per shape: milliseconds per kernel and the roofline figure of each
(correlate: N^2 g FLOP against the FP32-emulated tensor peak; the others: algorithmic bytes / s)."""
import argparse, ctypes, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from metacells_b200 import _lib

parser = argparse.ArgumentParser()
parser.add_argument("--cells", default="2000,4000,8000,16000,24000,32000,64000")
parser.add_argument("--genes", default="1000,2000,4000")
parser.add_argument("--repeat", type=int, default=3)
args = parser.parse_args()
library = _lib.lib()
for name in ("mcb_dev_correlate_dense_f32", "mcb_dev_rank_matrix_f32", "mcb_dev_collect_top_f32", "mcb_dev_collect_pruned_f32_i32_i32",
             "mcb_profile_enable", "mcb_profile_reset", "mcb_profile_read"):
    getattr(library, name).restype = ctypes.c_int
p = lambda tensor: ctypes.c_void_p(tensor.data_ptr())
i64 = ctypes.c_int64

def read(name):
    total, calls = ctypes.c_double(0), ctypes.c_longlong(0)
    library.mcb_profile_read(name.encode(), ctypes.byref(total), ctypes.byref(calls))
    return total.value / max(calls.value, 1)

peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
tensor_peak = peaks.get("bf16_tflops_sustained", 1400.0) / 6.0
print("| cells | genes | k | correlate ms | TFLOP/s (frac of %.0f) | rank_matrix ms | GB/s | collect_top ms | GB/s | collect_pruned ms | GB/s |" % tensor_peak)
print("|---|---|---|---|---|---|---|---|---|---|---|")
stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
for cells in [int(x) for x in args.cells.split(",")]:
    k = max(int(round(cells / 48)), 30)
    for genes in [int(x) for x in args.genes.split(",")]:
        generator = torch.Generator(device="cuda").manual_seed(123456)
        types = torch.randn((32, genes), device="cuda", generator=generator)
        member = torch.randint(0, 32, (cells,), device="cuda", generator=generator)
        dense = torch.log2(torch.poisson(torch.exp(1.0 + 0.6 * types[member]), generator=generator) + 1 / 7).contiguous()
        similarity = torch.empty((cells, cells), dtype=torch.float32, device="cuda")
        ranks = torch.empty_like(similarity)
        top_indices = torch.empty((cells, k), dtype=torch.int32, device="cuda")
        top_values = torch.empty((cells, k), dtype=torch.float32, device="cuda")
        indptr = (torch.arange(cells + 1, device="cuda", dtype=torch.int64) * k).to(torch.int32)
        pruned_degree = max(k // 3, 1)
        pruned_values = torch.empty(cells * pruned_degree, dtype=torch.float32, device="cuda")
        pruned_indices = torch.empty(cells * pruned_degree, dtype=torch.int32, device="cuda")
        pruned_indptr = torch.empty(cells + 1, dtype=torch.int32, device="cuda")
        library.mcb_profile_enable(ctypes.c_int(0))
        for attempt in range(args.repeat + 1):
            if attempt == 1:
                library.mcb_profile_reset()
                library.mcb_profile_enable(ctypes.c_int(1))
            _lib.check(library.mcb_dev_correlate_dense_f32(p(dense), i64(cells), i64(genes), i64(genes), p(similarity), i64(cells), ctypes.c_int(1), stream))
            ranks.copy_(similarity)
            _lib.check(library.mcb_dev_rank_matrix_f32(p(ranks), i64(cells), i64(cells), i64(cells), ctypes.c_int(0), stream))
            _lib.check(library.mcb_dev_collect_top_f32(i64(k), p(similarity), i64(cells), i64(cells), i64(cells), p(top_indices), p(top_values), ctypes.c_int(0), stream))
            _lib.check(library.mcb_dev_collect_pruned_f32_i32_i32(i64(pruned_degree), p(top_values), p(top_indices), p(indptr), i64(cells), p(pruned_values), p(pruned_indices), p(pruned_indptr), stream))
        torch.cuda.synchronize()
        correlate, rank, top, prune = read("correlate_umma"), read("rank_matrix") + read("rank_matrix_ties"), read("collect_top"), read("prune_bands")
        flops = float(cells) * cells * genes
        tf = flops / (correlate * 1e-3) / 1e12
        rank_gbs = 8.0 * cells * cells / (rank * 1e-3) / 1e9
        top_gbs = (4.0 * cells * cells + 8.0 * cells * k) / (top * 1e-3) / 1e9
        prune_gbs = (8.0 * cells * k + 8.0 * cells * pruned_degree) / (prune * 1e-3) / 1e9 if prune else 0.0
        print("| %d | %d | %d | %.3f | %.0f (%.2f) | %.3f | %.0f | %.3f | %.0f | %.3f | %.0f |" % (
            cells, genes, k, correlate, tf, tf / tensor_peak, rank, rank_gbs, top, top_gbs, prune, prune_gbs), flush=True)
        del dense, similarity, ranks
        torch.cuda.empty_cache()
