"""Developer diagnostic (GPU box): error of correlate_dense vs float64 Pearson for a few shapes."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from metacells_b200 import extensions as xt

def exact(d):
    with np.errstate(all="ignore"):
        e = np.atleast_2d(np.corrcoef(d.astype("float64")))
    e[np.isnan(e)] = 0
    np.fill_diagonal(e, 1)
    return e

rng = np.random.default_rng(0)
for shape in [(1, 5), (2, 3), (127, 31), (128, 32), (129, 33), (300, 1000), (513, 2049), (1500, 1000), (1000, 4000)]:
    d = rng.normal(-2.5, 0.6, size=shape).astype("float32")
    out = np.full((shape[0], shape[0]), 7.0, dtype="float32")
    xt.correlate_dense_float32_t(d, out)
    e = exact(d)
    err = np.abs(out - e)
    i, j = np.unravel_index(err.argmax(), err.shape)
    print(shape, "maxerr %.3g at (%d,%d) got %.6f want %.6f" % (err.max(), i, j, out[i, j], e[i, j]),
          "sym", np.array_equal(out, out.T), "n7", int((out == 7).sum()), flush=True)
