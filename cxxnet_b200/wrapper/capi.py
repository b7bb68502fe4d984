"""Python side of the C ABI (``libcxxnetwrapper.so``, ``include/cxxnet_wrapper.h``).

The shim in ``csrc/capi/cxxnet_wrapper.cpp`` embeds (or joins) a CPython interpreter and forwards
each of the 22 ``CXNIO*`` / ``CXNNet*`` functions of the reference (``wrapper/cxxnet_wrapper.h:29-225``)
to the functions below.  Handles are small integers; pointers cross the boundary as integers and
are read / written with ctypes.  As in the reference, returned pointers stay valid only until the
next call on the same handle (``cxxnet_wrapper.h:163-165``).
"""
from __future__ import annotations

import ctypes
from typing import Dict

import numpy as np

from .cxxnet import DataIter, Net

_objs: Dict[int, object] = {}
_keep: Dict[int, object] = {}
_next = [1]


def _new(o) -> int:
    h = _next[0]
    _next[0] += 1
    _objs[h] = o
    return h


def _arr(ptr: int, shape) -> np.ndarray:
    n = int(np.prod(shape))
    buf = (ctypes.c_float * n).from_address(ptr)
    return np.frombuffer(buf, dtype=np.float32).reshape(shape).copy()


def _uints(ptr: int, n: int):
    return (ctypes.c_uint * n).from_address(ptr)


def _ret(handle: int, a: np.ndarray) -> int:
    a = np.ascontiguousarray(a, dtype=np.float32)
    _keep[handle] = a
    return a.ctypes.data


# ---- iterators --------------------------------------------------------------------------------
def io_create(cfg: str) -> int:
    return _new(DataIter(cfg))


def io_next(h: int) -> int:
    return 1 if _objs[h].next() else 0


def io_before_first(h: int) -> int:
    _objs[h].before_first()
    return 0


def io_get_data(h: int, oshape: int, ostride: int) -> int:
    a = _objs[h].get_data()
    s = _uints(oshape, 4)
    for i in range(4):
        s[i] = a.shape[i]
    _uints(ostride, 1)[0] = a.shape[3]
    return _ret(h, a)


def io_get_label(h: int, oshape: int, ostride: int) -> int:
    a = _objs[h].get_label()
    s = _uints(oshape, 2)
    s[0], s[1] = a.shape[0], a.shape[1]
    _uints(ostride, 1)[0] = a.shape[1]
    return _ret(h, a)


def free(h: int) -> int:
    _objs.pop(h, None)
    _keep.pop(h, None)
    return 0


# ---- net --------------------------------------------------------------------------------------
def net_create(device, cfg: str) -> int:
    return _new(Net(dev=device or "", cfg=cfg or ""))


def net_set_param(h: int, name: str, val: str) -> int:
    _objs[h].set_param(name, val)
    return 0


def net_init_model(h: int) -> int:
    _objs[h].init_model()
    return 0


def net_save_model(h: int, fname: str) -> int:
    _objs[h].save_model(fname)
    return 0


def net_load_model(h: int, fname: str) -> int:
    _objs[h].load_model(fname)
    return 0


def net_start_round(h: int, r: int) -> int:
    _objs[h].start_round(r)
    return 0


def net_set_weight(h: int, p_weight: int, size: int, layer_name: str, tag: str) -> int:
    _objs[h].set_weight(_arr(p_weight, (size,)), layer_name, tag)
    return 0


def net_get_weight(h: int, layer_name: str, tag: str, wshape: int, out_dim: int) -> int:
    w = _objs[h].get_weight(layer_name, tag)
    if w is None:
        _uints(out_dim, 1)[0] = 0
        return 0
    s = _uints(wshape, 4)
    for i, d in enumerate(w.shape):
        s[i] = d
    _uints(out_dim, 1)[0] = w.ndim
    return _ret(h, w)


def net_update_iter(h: int, dh: int) -> int:
    _objs[h].update(_objs[dh])
    return 0


def net_update_batch(h: int, p_data: int, dshape: int, p_label: int, lshape: int) -> int:
    ds, ls = list(_uints(dshape, 4)), list(_uints(lshape, 2))
    _objs[h].update(_arr(p_data, ds), _arr(p_label, ls))
    return 0


def net_predict_batch(h: int, p_data: int, dshape: int, out_size: int) -> int:
    out = _objs[h].predict(_arr(p_data, list(_uints(dshape, 4))))
    _uints(out_size, 1)[0] = out.size
    return _ret(h, out)


def net_predict_iter(h: int, dh: int, out_size: int) -> int:
    out = _objs[h].predict(_objs[dh])
    _uints(out_size, 1)[0] = out.size
    return _ret(h, out)


def net_extract_batch(h: int, p_data: int, dshape: int, node_name: str, oshape: int) -> int:
    out = _objs[h].extract(_arr(p_data, list(_uints(dshape, 4))), node_name)
    s = _uints(oshape, 4)
    for i in range(4):
        s[i] = out.shape[i]
    return _ret(h, out)


def net_extract_iter(h: int, dh: int, node_name: str, oshape: int) -> int:
    out = _objs[h].extract(_objs[dh], node_name)
    s = _uints(oshape, 4)
    for i in range(4):
        s[i] = out.shape[i]
    return _ret(h, out)


def net_evaluate(h: int, dh: int, name: str) -> int:
    s = _objs[h].evaluate(_objs[dh], name).encode("utf-8") + b"\0"
    buf = ctypes.create_string_buffer(s)
    _keep[h] = buf
    return ctypes.addressof(buf)
