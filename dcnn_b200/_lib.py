"""ctypes binding of libdcnn_b200.so (the C ABI is declared in include/dcnn_b200.h).

The library is built in-tree (dcnn_b200/lib/) so it travels with the repository snapshot.  Loading is lazy and
loud: on a machine with a GPU a missing/unloadable library raises instead of silently falling back."""
from __future__ import annotations

import ctypes as C
import os

_LIB = None
LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "lib", "libdcnn_b200.so")


class NativeError(RuntimeError):
    pass


def lib():
    global _LIB
    if _LIB is None:
        path = os.environ.get("DCNN_B200_LIB", LIB_PATH)
        if not os.path.exists(path):
            from .build import build
            path = build()
        l = C.CDLL(path)
        l.dcnn_last_error.restype = C.c_char_p
        l.dcnn_version.restype = C.c_char_p
        l.dcnn_launch_count.restype = C.c_ulonglong
        _LIB = l
    return _LIB


def is_loaded() -> bool:
    return _LIB is not None


def check(rc: int, what: str = ""):
    if rc != 0:
        msg = lib().dcnn_last_error().decode(errors="replace")
        raise NativeError(f"{what} failed (code {rc}): {msg}")


def launch_count() -> int:
    return int(lib().dcnn_launch_count())


def reset_launch_count():
    lib().dcnn_reset_launch_count()


P = C.c_void_p
I = C.c_int
LL = C.c_longlong
ULL = C.c_ulonglong
F = C.c_float
D = C.c_double
