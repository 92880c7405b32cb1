"""ctypes binding of libdcnn_host.so (native CPU runtime: loaders, augmentation, codecs).  Built on first use."""
from __future__ import annotations

import ctypes as C
import os

_LIB = None
PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "lib", "libdcnn_host.so")


def lib():
    global _LIB
    if _LIB is None:
        path = PATH
        if not os.path.exists(path):
            from .build import build_host
            path = build_host()
        l = C.CDLL(path)
        l.dcnn_load_mnist_csv.restype = C.c_longlong
        l.dcnn_load_cifar_bin.restype = C.c_longlong
        l.dcnn_encode_message.restype = C.c_longlong
        l.dcnn_encode_tensor_record.restype = C.c_longlong
        _LIB = l
    return _LIB


def available() -> bool:
    try:
        lib()
        return True
    except Exception:
        return False
