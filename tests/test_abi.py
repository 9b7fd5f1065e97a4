"""CPU checks of the C-ABI boundary: the library loads and exports every symbol the header declares,
and the host layer refuses bad arguments before touching the GPU."""
import ctypes
import os

import numpy as np
import pytest

from metacells_b200 import _lib


def test_library_exports_every_declared_symbol():
    assert os.path.exists(_lib.LIB_PATH), "build libmcb200.so first: python -m metacells_b200.build"
    library = ctypes.CDLL(_lib.LIB_PATH)
    names = _lib.declared_symbols()
    assert len(names) >= 10
    missing = [name for name in names if not hasattr(library, name)]
    assert not missing, missing


def test_version_and_error_string():
    library = _lib.lib()
    library.mcb_version.restype = ctypes.c_int
    assert library.mcb_version() >= 1
    assert isinstance(library.mcb_last_error(), bytes)


def test_host_layer_rejects_wrong_dtype_and_shape():
    from metacells_b200 import extensions as xt

    dense = np.zeros((4, 8), dtype="float64")
    out = np.zeros((4, 4), dtype="float32")
    with pytest.raises(TypeError):
        xt.correlate_dense_float32_t(dense, out)
    with pytest.raises(ValueError):
        xt.correlate_dense_float64_t(dense, np.zeros((4, 5), dtype="float32"))
    with pytest.raises(ValueError):
        xt.correlate_dense_float64_t(dense[:, ::2], out)


def test_reference_symbol_names_exist():
    """Names computation.py builds at run time resolve on our module."""
    from metacells_b200 import extensions as xt

    for name in [
        "correlate_dense_float32_t", "correlate_dense_float64_t", "cross_correlate_dense_float32_t",
        "pairs_correlate_dense_float32_t", "downsample_compressed_float32_t_int32_t_float32_t",
        "downsample_dense_float32_t_float32_t", "downsample_array_float32_t_float32_t", "set_threads_count",
    ]:
        assert callable(getattr(xt, name)), name
