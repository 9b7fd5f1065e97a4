"""Row-block sharded similarity + top-k of an oversized pile over NCCL.  Launch with
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
      scripts/run_oversized_pile.py [--cells 40000] [--genes 1000] [--top 150]
Every rank computes its rows; rank 0 also computes the unsharded answer and checks that the gathered matrix equals it."""
import argparse, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch, torch.distributed as dist
from metacells_b200 import tools as tl

parser = argparse.ArgumentParser()
parser.add_argument("--cells", type=int, default=40000)
parser.add_argument("--genes", type=int, default=1000)
parser.add_argument("--top", type=int, default=150)
args = parser.parse_args()
rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
if world > 1:
    dist.init_process_group("nccl")
rng = np.random.default_rng(123456)
types = rng.normal(size=(16, args.genes)).astype("float32")
dense = np.log2(rng.poisson(np.exp(1.0 + 0.6 * types[rng.integers(0, 16, size=args.cells)])).astype("float32") + np.float32(1 / 7))
for attempt in range(3):
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    start = time.perf_counter()
    sharded = tl.compute_obs_obs_similarity_top(dense, args.top)
    torch.cuda.synchronize()
    elapsed = time.perf_counter() - start
    if rank == 0:
        print(f"world {world}: {args.cells} cells x {args.genes} genes, top {args.top}: {elapsed * 1e3:.1f} ms (host arrays in, CSR out)", flush=True)
if rank == 0:
    start = time.perf_counter()
    alone = tl.compute_obs_obs_similarity_top(dense, args.top, rank=0, world_size=1)
    print(f"unsharded on one GPU: {(time.perf_counter() - start) * 1e3:.1f} ms; identical: "
          f"{np.array_equal(alone.indices, sharded.indices) and np.array_equal(alone.data, sharded.data)}", flush=True)
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
