"""ctypes binding of the native data-pipeline library (``csrc/runtime``, ``libcxxnet_io.so``)."""
from __future__ import annotations

import ctypes
import os
from typing import List, Optional

_PATH = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "lib", "libcxxnet_io.so")
_lib = None


def lib():
    global _lib
    if _lib is None and os.path.exists(_PATH):
        L = ctypes.CDLL(_PATH)
        L.cxnio_reader_open.restype = ctypes.c_void_p
        L.cxnio_reader_open.argtypes = [ctypes.POINTER(ctypes.c_char_p), ctypes.c_int, ctypes.c_int]
        L.cxnio_reader_next.restype = ctypes.c_int
        L.cxnio_reader_next.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_void_p),
                                        ctypes.POINTER(ctypes.c_ulonglong)]
        L.cxnio_reader_rewind.argtypes = [ctypes.c_void_p]
        L.cxnio_reader_close.argtypes = [ctypes.c_void_p]
        L.cxnio_pack.restype = ctypes.c_longlong
        L.cxnio_pack.argtypes = [ctypes.c_char_p, ctypes.c_char_p, ctypes.c_char_p, ctypes.c_int]
        _lib = L
    return _lib


class PageReader:
    """blobs of one or more imgbin files, pages prefetched by a native thread."""

    def __init__(self, paths: List[str], nbuffers: int = 4):
        L = lib()
        if L is None:
            raise RuntimeError("libcxxnet_io.so is not built")
        arr = (ctypes.c_char_p * len(paths))(*[p.encode() for p in paths])
        self._h = L.cxnio_reader_open(arr, len(paths), nbuffers)

    def next(self) -> Optional[bytes]:
        p, n = ctypes.c_void_p(), ctypes.c_ulonglong()
        if not lib().cxnio_reader_next(self._h, ctypes.byref(p), ctypes.byref(n)):
            return None
        return ctypes.string_at(p.value, n.value)

    def rewind(self) -> None:
        lib().cxnio_reader_rewind(self._h)

    def close(self) -> None:
        if self._h:
            lib().cxnio_reader_close(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def pack(list_path: str, root: str, out_path: str, silent: bool = True) -> int:
    L = lib()
    if L is None:
        raise RuntimeError("libcxxnet_io.so is not built")
    return int(L.cxnio_pack(list_path.encode(), root.encode(), out_path.encode(), 1 if silent else 0))
