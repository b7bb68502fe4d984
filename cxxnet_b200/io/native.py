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
        if hasattr(L, "cxnio_reader_next_page"):
            L.cxnio_reader_next_page.restype = ctypes.c_int
            L.cxnio_reader_next_page.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        L.cxnio_reader_close.argtypes = [ctypes.c_void_p]
        L.cxnio_pack.restype = ctypes.c_longlong
        L.cxnio_pack.argtypes = [ctypes.c_char_p, ctypes.c_char_p, ctypes.c_char_p, ctypes.c_int]
        if hasattr(L, "cxnio_jpeg_info"):
            IP = ctypes.POINTER(ctypes.c_int)
            L.cxnio_jpeg_info.argtypes = [ctypes.c_void_p, ctypes.c_size_t, IP, IP, IP]
            L.cxnio_jpeg_decode_planar_f32.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p,
                                                       ctypes.c_int, ctypes.c_int]
            L.cxnio_jpeg_decode_rgb.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p,
                                                ctypes.c_int, ctypes.c_int]
            L.cxnio_jpeg_decode_many.argtypes = [ctypes.c_int, ctypes.POINTER(ctypes.c_void_p),
                                                 ctypes.POINTER(ctypes.c_size_t), ctypes.POINTER(ctypes.c_void_p),
                                                 IP, IP, IP, ctypes.c_int]
        if hasattr(L, "cxnio_image_info"):
            IP = ctypes.POINTER(ctypes.c_int)
            L.cxnio_image_info.argtypes = [ctypes.c_void_p, ctypes.c_size_t, IP, IP, IP]
            L.cxnio_image_decode_planar_f32.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p,
                                                        ctypes.c_int, ctypes.c_int]
            L.cxnio_raw_encode.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
            L.cxnio_raw_encode.restype = ctypes.c_size_t
            L.cxnio_pack_raw.restype = ctypes.c_longlong
            L.cxnio_pack_raw.argtypes = [ctypes.c_char_p, ctypes.c_char_p, ctypes.c_char_p, ctypes.c_int]
        if hasattr(L, "cxnio_decode_augment_batch"):
            L.cxnio_decode_augment_batch.argtypes = [
                ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_float, ctypes.c_void_p, ctypes.c_int,
                ctypes.POINTER(ctypes.c_int), ctypes.c_int]
            assert L.cxnio_aug_item_bytes() == ctypes.sizeof(AugItem)
        _lib = L
    return _lib


def _view(buf):
    """uint8 numpy view of a bytes-like object (no copy); `.ctypes.data` is its address"""
    import numpy as np
    return np.frombuffer(buf, dtype=np.uint8)


class AugItem(ctypes.Structure):
    """one instance of ``cxnio_decode_augment_batch`` (``csrc/runtime/batch_pipeline.cpp`` CxnioAugItem)"""
    _fields_ = [("buf", ctypes.c_void_p), ("size", ctypes.c_uint64), ("pixels", ctypes.c_void_p),
                ("src_h", ctypes.c_int32), ("src_w", ctypes.c_int32), ("y0", ctypes.c_int32), ("x0", ctypes.c_int32),
                ("flip", ctypes.c_int32), ("contrast", ctypes.c_float), ("illum", ctypes.c_float),
                ("affine", ctypes.c_int32), ("warp_h", ctypes.c_int32), ("warp_w", ctypes.c_int32),
                ("warp_y", ctypes.c_int32), ("warp_x", ctypes.c_int32), ("fill", ctypes.c_float),
                ("ai", ctypes.c_double * 6), ("interp", ctypes.c_int32), ("pad_", ctypes.c_int32)]


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

    def next_page(self, dst) -> bool:
        """the next whole page copied into `dst` (a writable uint8 numpy array of 64 MiB); False at the end"""
        assert dst.nbytes == 64 << 20
        return bool(lib().cxnio_reader_next_page(self._h, dst.ctypes.data))

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


def pack(list_path: str, root: str, out_path: str, silent: bool = True, raw: bool = False) -> int:
    """im2bin; raw = True decodes every (JPEG) file now and stores raw RGB records"""
    L = lib()
    if L is None:
        raise RuntimeError("libcxxnet_io.so is not built")
    if raw:
        return int(L.cxnio_pack_raw(list_path.encode(), root.encode(), out_path.encode(), 1 if silent else 0))
    return int(L.cxnio_pack(list_path.encode(), root.encode(), out_path.encode(), 1 if silent else 0))


def default_threads() -> int:
    """worker threads of a native batch call: this process's share of the cores (one process per GPU, so
    cores / ranks on this node), between 1 and 32"""
    cores = os.cpu_count() or 1
    try:
        cores = len(os.sched_getaffinity(0))
    except (AttributeError, OSError):
        pass
    ranks = max(int(os.environ.get("LOCAL_WORLD_SIZE", "1") or 1), 1)
    return max(1, min(32, cores // ranks))


# ---------------------------------------------------------------------------------------------- JPEG
def jpeg_info(buf: bytes):
    """(width, height, components) of a JPEG the native decoder can handle, else None."""
    L = lib()
    if L is None or not hasattr(L, "cxnio_jpeg_info") or buf[:2] != b"\xff\xd8":
        return None
    w, h, c = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    v = _view(buf)
    if L.cxnio_jpeg_info(v.ctypes.data, v.size, ctypes.byref(w), ctypes.byref(h), ctypes.byref(c)) != 0:
        return None
    return w.value, h.value, c.value


def image_info(buf: bytes):
    """(width, height, components) of a JPEG, a PNG or a raw record (``csrc/runtime/raw_image.cpp``) the native
    pipeline reads, else None"""
    L = lib()
    if L is None or not hasattr(L, "cxnio_image_info") or not (buf[:2] == b"\xff\xd8" or buf[:4] == b"CXNR"
                                                                    or buf[:4] == b"\x89PNG"):
        return None
    w, h, c = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    v = _view(buf)
    if L.cxnio_image_info(v.ctypes.data, v.size, ctypes.byref(w), ctypes.byref(h), ctypes.byref(c)) != 0:
        return None
    return w.value, h.value, c.value


def image_decode(buf: bytes):
    """JPEG or raw record -> float32 RGB planar (3, h, w), or None when the native pipeline does not read it"""
    import numpy as np
    info = image_info(buf)
    if info is None:
        return None
    w, h, _ = info
    out = np.empty((3, h, w), dtype=np.float32)
    v = _view(buf)
    if lib().cxnio_image_decode_planar_f32(v.ctypes.data, v.size, out.ctypes.data, w, h) != 0:
        return None
    return out


def raw_encode(rgb_hwc) -> bytes:
    """uint8 array (h, w, 3) -> raw record for an imgbin page"""
    import numpy as np
    a = np.ascontiguousarray(rgb_hwc, dtype=np.uint8)
    assert a.ndim == 3 and a.shape[2] == 3
    out = np.empty(12 + a.size, dtype=np.uint8)
    n = lib().cxnio_raw_encode(a.ctypes.data, a.shape[1], a.shape[0], out.ctypes.data)
    assert n == out.size
    return out.tobytes()


def jpeg_decode(buf: bytes):
    """JPEG bytes -> float32 RGB planar (3, h, w) with values 0..255 (``csrc/runtime/jpeg_decoder.cpp``);
    None when the file is not a JPEG this decoder supports (the caller then uses OpenCV / Pillow)."""
    import numpy as np
    info = jpeg_info(buf)
    if info is None:
        return None
    w, h, _ = info
    out = np.empty((3, h, w), dtype=np.float32)
    v = _view(buf)
    if lib().cxnio_jpeg_decode_planar_f32(v.ctypes.data, v.size, out.ctypes.data, w, h) != 0:
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
    views = [_view(bufs[i]) for i in idx]
    cb = (ctypes.c_void_p * n)(*[v.ctypes.data for v in views])
    cs = (ctypes.c_size_t * n)(*[len(bufs[i]) for i in idx])
    co = (ctypes.c_void_p * n)(*[a.ctypes.data for a in arrs])
    cw = (ctypes.c_int * n)(*[infos[i][0] for i in idx])
    ch = (ctypes.c_int * n)(*[infos[i][1] for i in idx])
    st = (ctypes.c_int * n)()
    lib().cxnio_jpeg_decode_many(n, cb, cs, co, cw, ch, st, nthreads or default_threads())
    for j, i in enumerate(idx):
        outs[i] = arrs[j] if st[j] == 0 else None
    return outs


def decode_augment_batch(items, shape, mean_mode: int, mean, scale: float, dst, dst_bf16: bool,
                         nthreads: int = 0) -> List[int]:
    """Decode, crop, mean-subtract, jitter, mirror and scale ``len(items)`` images into the rows of ``dst``
    (a contiguous torch tensor [n, C, H, W], C = 3 or 1, fp32 or bf16) on native worker threads.

    items: tuples ``(encoded bytes, decoded planar float array or None, src_h, src_w, y0, x0, flip, contrast,
    illum[, affine])`` where ``affine`` is None or ``(Ai, warp_h, warp_w, warp_y, warp_x, fill)`` as drawn by
    ``ImageAugmenter.draw`` (the crop is then taken from the warped image); ``mean_mode`` 0 none / 1 per-channel values / 2 mean image (per instance: subtracted before the
    crop when it has the decoded size, after it when it has the crop size).  Returns one status per item
    (0 = written)."""
    import numpy as np
    n = len(items)
    c, H, W = shape
    arr = (AugItem * n)()
    keep = []
    mean_h = mean_w = 0
    mean_ptr = None
    if mean is not None:
        mean = np.ascontiguousarray(mean, dtype=np.float32)
        mean_ptr = mean.ctypes.data
        if mean_mode == 2:
            mean_h, mean_w = mean.shape[1], mean.shape[2]
    status = (ctypes.c_int * n)()
    # a mean image is used in one of two ways depending on the instance's decoded size: split the batch
    groups = {}
    for i, item in enumerate(items):
        blob, pixels, h, w, y0, x0, flip, contrast, illum = item[:9]
        affine = item[9] if len(item) > 9 else None
        a = arr[i]
        if affine is not None:
            Ai, a.warp_h, a.warp_w, a.warp_y, a.warp_x, a.fill = affine[:6]
            a.interp = int(affine[6]) if len(affine) > 6 else 0
            a.affine = 1
            for k in range(6):
                a.ai[k] = float(Ai[k // 3, k % 3])
        if pixels is not None:
            pixels = np.ascontiguousarray(pixels, dtype=np.float32)
            keep.append(pixels)
            a.buf, a.size, a.pixels = None, 0, pixels.ctypes.data
        else:
            v = _view(blob)
            keep.append(v)
            a.buf, a.size, a.pixels = v.ctypes.data, v.size, None
        a.src_h, a.src_w, a.y0, a.x0, a.flip = h, w, y0, x0, 1 if flip else 0
        a.contrast, a.illum = contrast, illum
        mode = mean_mode
        eff = (H, W) if affine is not None else (h, w)      # after the affine stage the instance has the crop's size
        if mean_mode == 2 and eff != (mean_h, mean_w):
            mode = 3
        groups.setdefault(mode, []).append(i)
    assert dst.is_contiguous() and tuple(dst.shape[1:]) == (c, H, W) and dst.shape[0] >= n
    import torch
    kind = {torch.float32: 0, torch.bfloat16: 1, torch.uint8: 2}[dst.dtype]      # what the C side calls dst_bf16
    assert kind == (1 if dst_bf16 else kind) and (kind != 2 or (mean_mode == 0 and scale == 1.0))
    inst_bytes = c * H * W * dst.element_size()
    nthreads = nthreads or default_threads()
    for mode, idx in groups.items():
        # contiguous runs of instances that share a mode go down in one call
        run_start = 0
        while run_start < len(idx):
            run_end = run_start
            while run_end + 1 < len(idx) and idx[run_end + 1] == idx[run_end] + 1:
                run_end += 1
            i0, cnt = idx[run_start], run_end - run_start + 1
            lib().cxnio_decode_augment_batch(
                ctypes.addressof(arr) + i0 * ctypes.sizeof(AugItem), cnt, c, H, W, mode, mean_ptr,
                mean_h if mode == 2 else H, mean_w if mode == 2 else W, scale,
                dst.data_ptr() + i0 * inst_bytes, kind,
                ctypes.cast(ctypes.addressof(status) + 4 * i0, ctypes.POINTER(ctypes.c_int)), nthreads)
            run_start = run_end + 1
    return list(status)
