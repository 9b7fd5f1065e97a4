"""BASELINE.json config 5 on one B200: projection of a synthetic query onto a synthetic atlas (query x atlas Pearson
correlation + per-row ordered top-k), device resident, through the C-ABI (mcb_dev_project_top_f32).
  python scripts/run_projection.py [--query 100000] [--atlas 1000000] [--genes 2000] [--k 50] [--check 64]
Prints one JSON line: seconds for the whole projection (CUDA events, after one warm-up on a small slice), the share
and rate of the two kernels (correlate: 2 Nq Na g FLOP against the FP32-emulated tensor peak; ordered_top: 4 B per
correlation against the measured HBM bandwidth), and the parity of `--check` sampled query rows against a float64
torch restatement of correlate.cpp:433-494 + a stable descending argsort."""
import argparse, ctypes, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from metacells_b200 import _lib

parser = argparse.ArgumentParser()
parser.add_argument("--query", type=int, default=100000)
parser.add_argument("--atlas", type=int, default=1000000)
parser.add_argument("--genes", type=int, default=2000)
parser.add_argument("--k", type=int, default=50)
parser.add_argument("--check", type=int, default=64)
parser.add_argument("--types", type=int, default=200)
args = parser.parse_args()
library = _lib.lib()
p = lambda tensor: ctypes.c_void_p(tensor.data_ptr())
i64 = ctypes.c_int64
stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)

def read(name):
    total, calls = ctypes.c_double(0), ctypes.c_longlong(0)
    library.mcb_profile_read(name.encode(), ctypes.byref(total), ctypes.byref(calls))
    return total.value, calls.value

def synthetic(rows, generator, programs):
    """Log fractions of metacell-like rows: a mixture of two expression programs plus noise, in row chunks."""
    out = torch.empty((rows, args.genes), dtype=torch.float32, device="cuda")
    for first in range(0, rows, 65536):
        last = min(first + 65536, rows)
        a = torch.randint(0, args.types, (last - first,), device="cuda", generator=generator)
        b = torch.randint(0, args.types, (last - first,), device="cuda", generator=generator)
        w = torch.rand((last - first, 1), device="cuda", generator=generator)
        rate = torch.exp(w * programs[a] + (1 - w) * programs[b] + 0.15 * torch.randn((last - first, args.genes), device="cuda", generator=generator))
        out[first:last] = torch.log2(rate / rate.sum(dim=1, keepdim=True) + 1e-5)
    return out

generator = torch.Generator(device="cuda").manual_seed(123456)
programs = 1.2 * torch.randn((args.types, args.genes), device="cuda", generator=generator)
atlas = synthetic(args.atlas, generator, programs)
query = synthetic(args.query, generator, programs)
indices = torch.empty((args.query, args.k), dtype=torch.int32, device="cuda")
correlations = torch.empty((args.query, args.k), dtype=torch.float32, device="cuda")

def project(rows):
    _lib.check(library.mcb_dev_project_top_f32(p(query), i64(rows), p(atlas), i64(args.atlas), i64(args.genes), i64(args.k),
                                               p(indices), p(correlations), stream))

project(min(args.query, 256))  # warm-up: pools, function attributes
torch.cuda.synchronize()
library.mcb_profile_reset()
library.mcb_profile_enable(ctypes.c_int(1))
begin, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
begin.record()
project(args.query)
end.record()
torch.cuda.synchronize()
seconds = begin.elapsed_time(end) * 1e-3
library.mcb_profile_enable(ctypes.c_int(0))
umma_ms, umma_calls = read("correlate_umma")
top_ms, top_calls = read("ordered_top")
prepare_ms, _ = read("correlate_prepare")

# parity of sampled rows against float64
sample = torch.linspace(0, args.query - 1, min(args.check, args.query), device="cuda").long()
def centred(rows):
    rows = rows.double()
    rows = rows - rows.mean(dim=1, keepdim=True)
    return rows / rows.norm(dim=1, keepdim=True)
q64 = centred(query[sample])
worst_value, worst_order, mismatches = 0.0, 0.0, 0
exact = torch.empty((len(sample), args.atlas), dtype=torch.float64, device="cuda")
for first in range(0, args.atlas, 131072):
    last = min(first + 131072, args.atlas)
    exact[:, first:last] = q64 @ centred(atlas[first:last]).T
expected_values, expected_indices = torch.sort(exact, dim=1, descending=True, stable=True)
expected_values, expected_indices = expected_values[:, :args.k], expected_indices[:, :args.k]
mine_indices = indices[sample].long()
mine_values = correlations[sample].double()
worst_value = float((mine_values - torch.gather(exact, 1, mine_indices)).abs().max())
worst_order = float((torch.gather(exact, 1, mine_indices) - expected_values).abs().max())
mismatches = int((mine_indices != expected_indices).sum())

peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
tensor_peak = peaks.get("bf16_tflops_sustained", 1400.0) / 6.0
hbm = peaks.get("hbm_gbs", 6500.0)
flops = 2.0 * args.query * args.atlas * args.genes
top_bytes = 4.0 * args.query * args.atlas + 8.0 * args.query * args.k
print(json.dumps({
    "workload": f"projection: {args.query} query rows x {args.atlas} atlas rows x {args.genes} genes, top {args.k}",
    "seconds": seconds, "query_rows_per_s": args.query / seconds,
    "correlate_umma": {"ms": umma_ms, "launches": umma_calls, "tflops": flops / (umma_ms * 1e-3) / 1e12,
                       "frac_of_emulated_fp32_peak": flops / (umma_ms * 1e-3) / 1e12 / tensor_peak},
    "ordered_top": {"ms": top_ms, "launches": top_calls, "gbs": top_bytes / (top_ms * 1e-3) / 1e9,
                    "frac_of_hbm": top_bytes / (top_ms * 1e-3) / 1e9 / hbm},
    "correlate_prepare_ms": prepare_ms,
    "parity": {"rows_checked": int(len(sample)), "max_abs_correlation_error": worst_value,
               "max_abs_gap_to_exact_order": worst_order, "positions_with_another_atlas_row": mismatches,
               "tolerance": 1e-5},
}))
