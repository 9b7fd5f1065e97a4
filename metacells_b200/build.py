"""In-tree build of libmcb200.so (nvcc, sm_100a only).  Run: python -m metacells_b200.build"""
from __future__ import annotations

import glob
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "csrc", "build")
LIB = os.path.join(HERE, "libmcb200.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-ccbin", "/usr/bin/g++",
]


def _stale(target: str, sources) -> bool:
    if not os.path.exists(target):
        return True
    stamp = os.path.getmtime(target)
    return any(os.path.getmtime(source) > stamp for source in sources)


def build(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(OBJ, exist_ok=True)
    sources = sorted(glob.glob(os.path.join(CSRC, "*.cu")))
    headers = glob.glob(os.path.join(CSRC, "*.cuh")) + glob.glob(os.path.join(HERE, "..", "include", "*.h"))
    objects = []
    jobs = []
    for source in sources:
        obj = os.path.join(OBJ, os.path.basename(source)[:-3] + ".o")
        objects.append(obj)
        if force or _stale(obj, [source] + headers):
            jobs.append([NVCC] + FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", source, "-o", obj])

    def run(command):
        result = subprocess.run(command, capture_output=True, text=True)
        if result.returncode != 0 or verbose:
            sys.stderr.write(result.stdout + result.stderr)
        if result.returncode != 0:
            raise RuntimeError("nvcc failed: " + " ".join(command))

    with ThreadPoolExecutor(max_workers=8) as pool:
        list(pool.map(run, jobs))
    if jobs or force or _stale(LIB, objects):
        run([NVCC, "-shared", "-o", LIB] + objects + ["-ccbin", "/usr/bin/g++"])
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
