"""Developer diagnostic (GPU box): what the host -> device link gives for one pile's arrays (page-locked by torch, as
bench.py holds them), alone and under a kernel that streams HBM, one copy or the 16 chunk copies of mcb_pile_create_host.
  python scripts/time_upload.py [--mbytes 233]"""
import argparse, json, time
import torch

parser = argparse.ArgumentParser()
parser.add_argument("--mbytes", type=int, default=233)
args = parser.parse_args()
count = args.mbytes * (1 << 20) // 4
host = [torch.empty(count, dtype=torch.float32).pin_memory() for _ in range(2)]
for tensor in host:
    tensor.fill_(1.0)
device = [torch.empty(count, dtype=torch.float32, device="cuda") for _ in range(2)]
busy = torch.empty(1 << 28, dtype=torch.float32, device="cuda")
copy_stream = torch.cuda.Stream()
report = {"mbytes_per_array": args.mbytes}


def timed(name, chunks, under_kernels, repeat=5):
    best = None
    for _ in range(repeat):
        torch.cuda.synchronize()
        start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if under_kernels:
            for _ in range(12):
                busy.mul_(1.0001)
        with torch.cuda.stream(copy_stream):
            start.record()
            step = (count + chunks - 1) // chunks
            for chunk in range(chunks):
                low, high = chunk * step, min(count, (chunk + 1) * step)
                for source, target in zip(host, device):
                    target[low:high].copy_(source[low:high], non_blocking=True)
            end.record()
        torch.cuda.synchronize()
        ms = start.elapsed_time(end)
        best = ms if best is None else min(best, ms)
    report[name] = {"ms": round(best, 3), "GBps": round(2 * count * 4 / best / 1e6, 1)}


timed("one_copy_alone", 1, False)
timed("eight_chunks_alone", 8, False)
timed("one_copy_under_kernels", 1, True)
timed("eight_chunks_under_kernels", 8, True)
start = time.perf_counter()
torch.cuda.synchronize()
print(json.dumps(report))
