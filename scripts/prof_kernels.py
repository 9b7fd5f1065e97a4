"""Developer profiling driver (GPU box): one small launch of each non-tensor kernel of the path."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
from synth import synth_umis
from metacells_b200 import extensions as xt, tools as tl

which = sys.argv[1] if len(sys.argv) > 1 else "all"
rng = np.random.default_rng(0)
if which in ("all", "rank"):
    sim = np.abs(rng.normal(size=(592, 16000))).astype("float32")
    sim /= sim.max()
    xt.rank_matrix_float32_t(sim, False)
if which in ("all", "downsample"):
    umis = synth_umis(1184, 27000, seed=1)
    out = np.empty_like(umis.data)
    xt.downsample_compressed_float32_t_int32_t_float32_t(umis.data, umis.indptr, out, np.full(1184, 3296, dtype="int32"), 123456)
if which in ("all", "knn"):
    base = rng.normal(size=(4096, 16)).astype("float32")
    sim = np.abs(np.corrcoef(base)).astype("float32")
    np.fill_diagonal(sim, 1.0)
    tl.compute_obs_obs_knn_graph(sim, k=85)
print("done")
