"""Drop-in shim: makes ``import metacells.extensions`` resolve to the B200 library for the hot-path symbols.

Put this directory's parent (``dropin/``) on ``sys.path`` BEFORE the installed ``metacells`` package (and the repository
root, for ``metacells_b200``).  ``import metacells`` then finds this file first.  It locates the stock package further
down ``sys.path``, makes it part of this package's ``__path__`` (so ``metacells.utilities``, ``metacells.tools``, ... are
the stock modules, unchanged) and runs the stock ``__init__.py`` in this module's namespace; the one module that comes
from here is ``metacells.extensions`` (extensions.py beside this file), which the stock Python code imports as ``xt``
(utilities/computation.py:53, utilities/parallel.py:76, tools/candidates.py:23).  The reference registers that module
at metacells/extensions.cpp:64-83.
"""
import importlib.machinery as _machinery
import os as _os
import sys as _sys

_HERE = _os.path.dirname(_os.path.abspath(__file__))


def _find_stock_package():
    for entry in _sys.path:
        candidate = _os.path.join(entry or ".", "metacells")
        if _os.path.abspath(candidate) == _HERE:
            continue
        if _os.path.isfile(_os.path.join(candidate, "__init__.py")):
            return _os.path.abspath(candidate)
    return None


STOCK_PACKAGE_DIRECTORY = _find_stock_package()
if STOCK_PACKAGE_DIRECTORY is not None:
    __path__ = [_HERE, STOCK_PACKAGE_DIRECTORY]  # extensions.py from here, everything else from the stock package
    with open(_os.path.join(STOCK_PACKAGE_DIRECTORY, "__init__.py"), "r", encoding="utf-8") as _handle:
        exec(compile(_handle.read(), _os.path.join(STOCK_PACKAGE_DIRECTORY, "__init__.py"), "exec"), globals())  # noqa: S102


def stock_extension_path():
    """The stock compiled ``extensions`` module (a pybind11 shared object) of the installed package, or None."""
    if STOCK_PACKAGE_DIRECTORY is None:
        return None
    for suffix in _machinery.EXTENSION_SUFFIXES:
        candidate = _os.path.join(STOCK_PACKAGE_DIRECTORY, "extensions" + suffix)
        if _os.path.isfile(candidate):
            return candidate
    return None
