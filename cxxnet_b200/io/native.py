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
        if hasattr(L, "cxnio_jpeg_info"):
            IP = ctypes.POINTER(ctypes.c_int)
            L.cxnio_jpeg_info.argtypes = [ctypes.c_char_p, ctypes.c_size_t, IP, IP, IP]
            L.cxnio_jpeg_decode_planar_f32.argtypes = [ctypes.c_char_p, ctypes.c_size_t, ctypes.c_void_p,
                                                       ctypes.c_int, ctypes.c_int]
            L.cxnio_jpeg_decode_rgb.argtypes = [ctypes.c_char_p, ctypes.c_size_t, ctypes.c_void_p,
                                                ctypes.c_int, ctypes.c_int]
            L.cxnio_jpeg_decode_many.argtypes = [ctypes.c_int, ctypes.POINTER(ctypes.c_char_p),
                                                 ctypes.POINTER(ctypes.c_size_t), ctypes.POINTER(ctypes.c_void_p),
                                                 IP, IP, IP, ctypes.c_int]
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


# ---------------------------------------------------------------------------------------------- JPEG
def jpeg_info(buf: bytes):
    """(width, height, components) of a JPEG the native decoder can handle, else None."""
    L = lib()
    if L is None or not hasattr(L, "cxnio_jpeg_info") or buf[:2] != b"\xff\xd8":
        return None
    w, h, c = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    if L.cxnio_jpeg_info(buf, len(buf), ctypes.byref(w), ctypes.byref(h), ctypes.byref(c)) != 0:
        return None
    return w.value, h.value, c.value


def jpeg_decode(buf: bytes):
    """JPEG bytes -> float32 RGB planar (3, h, w) with values 0..255 (``csrc/runtime/jpeg_decoder.cpp``);
    None when the file is not a JPEG this decoder supports (the caller then uses OpenCV / Pillow)."""
    import numpy as np
    info = jpeg_info(buf)
    if info is None:
        return None
    w, h, _ = info
    out = np.empty((3, h, w), dtype=np.float32)
    if lib().cxnio_jpeg_decode_planar_f32(buf, len(buf), out.ctypes.data, w, h) != 0:
        return None
    return out


def jpeg_decode_many(bufs: List[bytes], nthreads: int = 0) -> list:
    """decode a list of JPEGs on ``nthreads`` native worker threads (the GIL is released for the whole call);
    entries the decoder cannot handle come back as None."""
    import numpy as np
    infos = [jpeg_info(b) for b in bufs]
    idx = [i for i, inf in enumerate(infos) if inf is not None]
    outs: list = [None] * len(bufs)
    if not idx:
        return outs
    n = len(idx)
    arrs = [np.empty((3, infos[i][1], infos[i][0]), dtype=np.float32) for i in idx]
    cb = (ctypes.c_char_p * n)(*[bufs[i] for i in idx])
    cs = (ctypes.c_size_t * n)(*[len(bufs[i]) for i in idx])
    co = (ctypes.c_void_p * n)(*[a.ctypes.data for a in arrs])
    cw = (ctypes.c_int * n)(*[infos[i][0] for i in idx])
    ch = (ctypes.c_int * n)(*[infos[i][1] for i in idx])
    st = (ctypes.c_int * n)()
    lib().cxnio_jpeg_decode_many(n, cb, cs, co, cw, ch, st, nthreads or min(8, os.cpu_count() or 1))
    for j, i in enumerate(idx):
        outs[i] = arrs[j] if st[j] == 0 else None
    return outs
