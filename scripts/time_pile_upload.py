"""Developer diagnostic (GPU box): mcb_pile_create_host + mcb_pile_wait_upload of the bench pile from page-locked memory,
alone, under a kernel that streams HBM, and with one copy per array instead of eight chunks.
  python scripts/time_pile_upload.py"""
import ctypes, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
from synth import synth_umis
from metacells_b200 import _lib

umis = synth_umis(16000, 27000, seed=123456)
pinned = {name: torch.from_numpy(array).pin_memory() for name, array in
          (("data", umis.data.astype("float32")), ("indices", umis.indices.astype("int32")), ("indptr", umis.indptr.astype("int32")))}
p = ctypes.c_void_p
busy = torch.empty(1 << 28, dtype=torch.float32, device="cuda")
report = {"bytes": int(umis.nnz) * 8}


def upload_once():
    pile = ctypes.c_void_p()
    library = _lib.lib()
    _lib.check(library.mcb_pile_create_host(p(pinned["data"].data_ptr()), p(pinned["indices"].data_ptr()), p(pinned["indptr"].data_ptr()),
                                            ctypes.c_int64(16000), ctypes.c_int64(27000), ctypes.byref(pile)))
    _lib.check(library.mcb_pile_wait_upload(pile))
    return pile


def timed(name, under_kernels, repeat=6):
    best = None
    for _ in range(repeat):
        torch.cuda.synchronize()
        if under_kernels:
            for _ in range(16):
                busy.mul_(1.0001)
        start = time.perf_counter()
        pile = upload_once()
        ms = (time.perf_counter() - start) * 1e3
        torch.cuda.synchronize()
        _lib.lib().mcb_pile_free(pile)
        best = ms if best is None else min(best, ms)
    report[name] = {"ms": round(best, 3), "GBps": round(report["bytes"] / best / 1e6, 1)}


timed("eight_chunks_alone", False)
timed("eight_chunks_under_kernels", True)
os.environ["MCB_PILE_CHUNKS"] = "1"
timed("one_chunk_alone", False)
timed("one_chunk_under_kernels", True)
os.environ["MCB_PILE_CHUNKS"] = "32"
timed("32_chunks_alone", False)
print(json.dumps(report))
