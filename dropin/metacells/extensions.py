"""``metacells.extensions`` for a B200: the hot-path symbols come from ``metacells_b200.extensions``
(libmcb200.so through the C-ABI of include/mcb200.h); every other symbol of the reference's 2 506 exports falls through to
the stock compiled module of the installed package, loaded lazily on first use.  Python builds the names at run time
(``getattr(xt, f"correlate_dense_{dtype}_t")``, utilities/computation.py:461, :879-883, :1374-1378, :1700), so both
kinds are served by ``__getattr__`` / plain attributes of this module."""
import importlib.util as _util
import threading as _threading

from metacells_b200 import extensions as _b200

#: the names served by the GPU library
HOT_PATH_SYMBOLS = tuple(sorted(
    name for name in dir(_b200)
    if not name.startswith("_") and callable(getattr(_b200, name)) and name not in ("McbError", "call", "call_f64", "host_ptr",
                                                                                  "row_stride", "glibc_log")))
for _name in HOT_PATH_SYMBOLS:
    globals()[_name] = getattr(_b200, _name)

_stock = None
_stock_lock = _threading.Lock()


def _stock_module():
    global _stock
    with _stock_lock:
        if _stock is None:
            import metacells

            path = metacells.stock_extension_path()
            if path is None:
                raise AttributeError("no stock metacells.extensions module is installed to fall through to")
            spec = _util.spec_from_file_location("metacells_stock.extensions", path)  # PyInit_extensions
            module = _util.module_from_spec(spec)
            spec.loader.exec_module(module)
            _stock = module
    return _stock


def set_threads_count(count):
    """extensions.cpp:13-18: the GPU path has no host thread pool; the stock module (when present) still gets the count
    for the functions that fall through to it."""
    _b200.set_threads_count(count)
    try:
        _stock_module().set_threads_count(count)
    except AttributeError:
        pass


def __getattr__(name):
    if name.startswith("__"):
        raise AttributeError(name)
    return getattr(_stock_module(), name)
