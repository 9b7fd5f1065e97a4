"""Loader of libmcb200.so (the C-ABI of include/mcb200.h).  There is no CPU fallback: if the
library is missing, or an entry point reports an error, the caller gets an exception."""
from __future__ import annotations

import ctypes
import os
import re
from typing import Dict, List

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libmcb200.so")
HEADER_PATH = os.path.join(HERE, "..", "include", "mcb200.h")

_lib = None


class McbError(RuntimeError):
    """An entry point of libmcb200.so returned a non-zero code."""


def declared_symbols() -> List[str]:
    """Every function name include/mcb200.h declares."""
    with open(HEADER_PATH, "r", encoding="utf-8") as handle:
        text = handle.read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(mcb_[a-z0-9_]+)\s*\(", text)))


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise McbError(
                f"{LIB_PATH} is missing: build it with `python -m metacells_b200.build` "
                "(there is no CPU fallback for the metacells_b200 hot path)")
        _lib = ctypes.CDLL(LIB_PATH)
        _lib.mcb_last_error.restype = ctypes.c_char_p
    return _lib


def check(code: int) -> None:
    if code != 0:
        raise McbError(f"libmcb200 error {code}: {lib().mcb_last_error().decode('utf-8', 'replace')}")


_CTYPES: Dict[str, object] = {
    "i64": ctypes.c_int64,
    "u64": ctypes.c_uint64,
    "i32": ctypes.c_int32,
    "u32": ctypes.c_uint32,
    "int": ctypes.c_int,
    "f32": ctypes.c_float,
    "f64": ctypes.c_double,
    "ptr": ctypes.c_void_p,
}


def call(name: str, *args) -> None:
    """Call `name` with (kind, value) pairs; kind is a key of _CTYPES."""
    function = getattr(lib(), name)
    converted = [_CTYPES[kind](value) for kind, value in args]
    function.restype = ctypes.c_int
    check(function(*converted))


def call_f64(name: str, *args) -> float:
    """As `call`, for entry points whose last parameter is a `double*` result."""
    result = ctypes.c_double(0.0)
    function = getattr(lib(), name)
    converted = [_CTYPES[kind](value) for kind, value in args] + [ctypes.byref(result)]
    function.restype = ctypes.c_int
    check(function(*converted))
    return result.value


def host_ptr(array: np.ndarray, dtype: str, name: str, writable: bool = False):
    """('ptr', address) of a numpy array after the checks the reference's views make
    (extensions.h:141-235: exact dtype, unit element stride)."""
    if not isinstance(array, np.ndarray):
        raise TypeError(f"{name}: expected a numpy array, got {type(array).__name__}")
    if array.dtype != np.dtype(dtype):
        raise TypeError(f"{name}: expected dtype {dtype}, got {array.dtype}")
    if writable and not array.flags.writeable:
        raise ValueError(f"{name}: output array is read-only")
    if array.ndim == 1 and array.size > 1 and array.strides[0] != array.itemsize:
        raise ValueError(f"{name}: vector must have unit stride")
    if array.ndim == 2 and array.shape[1] > 1 and array.strides[1] != array.itemsize:
        raise ValueError(f"{name}: matrix rows must be contiguous")
    return ("ptr", array.ctypes.data)


def row_stride(matrix: np.ndarray) -> int:
    """Row stride in elements of a row-major (possibly row-sliced) matrix (extensions.h:270-277)."""
    if matrix.shape[0] <= 1:
        return max(matrix.shape[1], 1)
    stride = matrix.strides[0] // matrix.itemsize
    if matrix.strides[0] % matrix.itemsize != 0 or stride < matrix.shape[1]:
        raise ValueError("matrix row stride must be a multiple of the element size and >= columns")
    return stride
