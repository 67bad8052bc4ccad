"""Loader of the product library libczb200_{f32,f64}.so (built in-tree by
cubez_b200/csrc/Makefile).  There is NO fallback: a missing library is an ImportError-like
RuntimeError, and the library itself aborts when no CUDA device is present."""
import ctypes as C
import os

from . import ffi

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIBS = {}


def lib_path(real="f32"):
    return os.path.join(_HERE, "libczb200_%s.so" % real)


def load(real="f32"):
    """Load (once) and return the ctypes handle of libczb200_<real>.so with signatures bound."""
    if real not in ("f32", "f64"):
        raise ValueError("real must be 'f32' or 'f64'")
    if real in _LIBS:
        return _LIBS[real]
    path = lib_path(real)
    if not os.path.exists(path):
        raise RuntimeError(
            "cubez_b200: %s is missing -- build it with `make -C cubez_b200/csrc` "
            "(or __graft_entry__.build()). There is no CPU/PyTorch fallback." % path)
    dll = C.CDLL(path, mode=C.RTLD_LOCAL)
    ffi.bind_reference_symbols(dll)
    ffi.bind_czb_symbols(dll, C.c_float if real == "f32" else C.c_double)
    assert dll.czb_real_bytes() == (4 if real == "f32" else 8)
    _LIBS[real] = dll
    return dll
