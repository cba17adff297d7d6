"""ctypes binding of libeksb.so (the sm_100a CUDA library behind this package).

There is NO CPU fallback: if the shared library is missing or fails to load, importing
anything that computes raises, loudly.  Build it with ``make -C eks_b200/csrc -j8`` or
``python -c "import __graft_entry__ as g; g.build()"``.

Prototypes are read from ``include/eksb.h`` so the Python side can never drift from the
C ABI the header declares.
"""
from __future__ import annotations

import ctypes
import os
import re

_HERE = os.path.dirname(os.path.abspath(__file__))
# EKSB_LIBRARY: developer override to load a tuning build of the same library (tools/tune.sh)
LIB_PATH = os.environ.get("EKSB_LIBRARY") or os.path.join(_HERE, "libeksb.so")
HEADER_PATH = os.path.join(os.path.dirname(_HERE), "include", "eksb.h")

_CTYPES = {
    "int": ctypes.c_int, "long": ctypes.c_long, "double": ctypes.c_double, "float": ctypes.c_float,
    "void*": ctypes.c_void_p, "int*": ctypes.c_void_p, "double*": ctypes.c_void_p, "float*": ctypes.c_void_p,
    "const double*": ctypes.c_void_p, "const float*": ctypes.c_void_p, "const int*": ctypes.c_void_p,
    "const char*": ctypes.c_char_p, "size_t": ctypes.c_size_t,
    "const long*": ctypes.c_void_p, "long*": ctypes.c_void_p, "char*": ctypes.c_void_p,
}


class EksbError(RuntimeError):
    """A libeksb call returned a negative status."""


def parse_header(path=HEADER_PATH):
    """-> {name: (restype_str, [(type_str, arg_name), ...])} for every EKSB_API declaration."""
    text = open(path).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    protos = {}
    for m in re.finditer(r"EKSB_API\s+([\w\s\*]+?)\s*(eksb_\w+)\s*\(([^)]*)\)\s*;", text):
        ret, name, args = m.group(1).strip(), m.group(2), m.group(3).strip()
        arglist = []
        if args and args != "void":
            for a in args.split(","):
                a = " ".join(a.split())
                mm = re.match(r"(.*?)(\w+)$", a)
                ty = mm.group(1).strip().replace(" *", "*")
                arglist.append((ty, mm.group(2)))
        protos[name] = (ret.replace(" *", "*"), arglist)
    return protos


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"eks_b200: {LIB_PATH} not found. This package has no CPU fallback; build the CUDA "
            "library first (make -C eks_b200/csrc -j8).")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (ret, args) in parse_header().items():
        fn = getattr(lib, name)          # AttributeError here = header/library mismatch
        fn.restype = _CTYPES[ret]
        fn.argtypes = [_CTYPES[t] for t, _ in args]
    return lib


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = _load()
    return _lib


def check(rc, what):
    if rc != 0:
        msg = lib().eksb_last_error_string().decode()
        raise EksbError(f"{what} failed with status {rc}: {msg}")


# constants mirrored from include/eksb.h
LANES = 32
SHARED_M0, SHARED_S0, SHARED_C, SHARED_A, SHARED_Q, SHARED_ALL = 1, 2, 4, 8, 16, 31
A_DIAG, A_DENSE = 0, 1
MODE_MEDIAN, MODE_MEAN = 0, 1
FIRST_PRIOR_UPDATE, FIRST_PRIOR_ONLY = 1, 2
STATUS_PIVOT, STATUS_NONPOS_VAR = 1, 2
SUPPORTED_DIMS = ((2, 4), (3, 4), (3, 6), (3, 8), (4, 6), (6, 8))
