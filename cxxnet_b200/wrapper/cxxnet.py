"""Drop-in replacement for the reference's ``wrapper/cxxnet.py`` (Python 3).

Same surface (reference ``wrapper/cxxnet.py:64-312``): ``DataIter{next, before_first, check_valid,
get_data, get_label}``, ``Net{set_param, init_model, load_model, save_model, start_round, update,
evaluate, predict, extract, set_weight, get_weight}`` and ``train``.  The reference binds a C ABI
through ctypes; here the objects drive ``cxxnet_b200.trainer.NetTrainer`` directly (the same
C ABI is also available to non-Python callers as ``libcxxnetwrapper.so``, see ``capi.py``).

Reference quirks that are fixed rather than reproduced: the ``dev``
argument is honoured (the reference overwrites it with "gpu"), ``start_round`` reaches the
trainer, 1-D labels are reshaped with ``label.shape[0]``, and both ``train`` signatures work.
"""
from __future__ import annotations

import struct
import sys
from typing import Optional

import numpy as np
import torch

from ..config import parse_config_string
from ..io.data import DataBatch, create_iterator
from ..trainer import NetTrainer


class DataIter:
    """data iterator of cxxnet, built from a config string (an ``iter = ... iter = end`` block)."""

    def __init__(self, cfg: str):
        pairs = parse_config_string(cfg)
        self._it = create_iterator(pairs)
        self._it.init()
        self._it.before_first()
        self.head = True
        self.tail = False

    def __del__(self):
        try:
            self._it.close()
        except Exception:
            pass

    def next(self) -> bool:
        ret = self._it.next()
        self.head = False
        self.tail = not ret
        return bool(ret)

    def before_first(self) -> None:
        self._it.before_first()
        self.head = True
        self.tail = False

    def check_valid(self) -> None:
        if self.head:
            raise Exception("iterator was at head state, call next to get to valid state")
        if self.tail:
            raise Exception("iterator reaches end")

    def value(self) -> DataBatch:
        self.check_valid()
        return self._it.value()

    def get_data(self) -> np.ndarray:
        return self.value().data.numpy().astype(np.float32, copy=True)

    def get_label(self) -> np.ndarray:
        return self.value().label.numpy().astype(np.float32, copy=True)


def _batch_from_arrays(data: np.ndarray, label: Optional[np.ndarray]) -> DataBatch:
    b = DataBatch()
    b.data = torch.from_numpy(np.ascontiguousarray(data, dtype=np.float32))
    n = data.shape[0]
    if label is None:
        b.label = torch.zeros((n, 1), dtype=torch.float32)
    else:
        b.label = torch.from_numpy(np.ascontiguousarray(label, dtype=np.float32))
    b.batch_size = n
    b.inst_index = np.arange(n, dtype=np.uint32)
    return b


class Net:
    """neural net object"""

    def __init__(self, dev: str = "cpu", cfg: str = ""):
        self._tr = NetTrainer()
        self._cfg = parse_config_string(cfg)
        self._dev = dev
        self._net_type = 0
        self._round = 0
        self._inited = False
        for k, v in self._cfg:
            self._tr.set_param(k, v)
        if dev:
            self._tr.set_param("dev", dev)

    def set_param(self, name, value) -> None:
        name, value = str(name), str(value)
        if self._inited:
            raise RuntimeError("Net.set_param: parameters must be set before init_model / load_model")
        self._tr.set_param(name, value)

    def init_model(self) -> None:
        self._tr.init_model()
        self._inited = True

    def load_model(self, fname: str) -> None:
        with open(fname, "rb") as f:
            self._net_type = struct.unpack("<i", f.read(4))[0]
            self._tr.load_model(f)
        self._inited = True

    def save_model(self, fname: str) -> None:
        with open(fname, "wb") as f:
            f.write(struct.pack("<i", self._net_type))
            self._tr.save_model(f)

    def start_round(self, round_counter: int) -> None:
        self._round = int(round_counter)
        self._tr.start_round(self._round)

    def _check4(self, data, who):
        if data.ndim != 4:
            raise Exception("%s: need 4 dimensional tensor (batch, channel, height, width)" % who)

    def update(self, data, label=None) -> None:
        if isinstance(data, DataIter):
            self._tr.update(data.value())
            return
        if isinstance(data, np.ndarray):
            self._check4(data, "Net.update")
            if label is None:
                raise Exception("Net.update: need label to use update")
            if not isinstance(label, np.ndarray):
                raise Exception("Net.update: label need to be ndarray")
            if label.ndim == 1:
                label = label.reshape((label.shape[0], 1))
            if label.ndim != 2:
                raise Exception("Net.update: label need to be 2 dimension or one dimension ndarray")
            if label.shape[0] != data.shape[0]:
                raise Exception("Net.update: data size mismatch")
            self._tr.update(_batch_from_arrays(data, label))
            return
        raise Exception("update do not support type %s" % str(type(data)))

    def evaluate(self, data, name: str) -> str:
        if isinstance(data, DataIter):
            return self._tr.evaluate(data._it, name)
        raise Exception("update do not support type %s" % str(type(data)))

    def predict(self, data) -> np.ndarray:
        if isinstance(data, DataIter):
            b = data.value()
            out = self._tr.predict(b)
            return out[: b.batch_size - b.num_batch_padd] if b.num_batch_padd else out
        if isinstance(data, np.ndarray):
            self._check4(data, "Net.predict")
            return self._tr.predict(_batch_from_arrays(data, None))
        raise Exception("predict do not support type %s" % str(type(data)))

    def extract(self, data, name: str) -> np.ndarray:
        if isinstance(data, DataIter):
            return self._tr.extract_feature(data.value(), name)
        if isinstance(data, np.ndarray):
            self._check4(data, "Net.extract")
            return self._tr.extract_feature(_batch_from_arrays(data, None), name)
        raise Exception("extract do not support type %s" % str(type(data)))

    def set_weight(self, weight: np.ndarray, layer_name: str, tag: str) -> None:
        if tag not in ("bias", "wmat"):
            raise Exception("tag must be bias or wmat")
        self._tr.set_weight(np.ascontiguousarray(weight, dtype=np.float32), layer_name, tag)

    def get_weight(self, layer_name: str, tag: str):
        if tag not in ("bias", "wmat"):
            raise Exception("tag must be bias or wmat")
        w, shape = self._tr.get_weight(layer_name, tag)
        if w is None:
            return None
        return w.reshape(shape).astype(np.float32)


def train(cfg, data, *args, **kw):
    """``train(cfg, data_iter, num_round, param, eval_data=None)`` or
    ``train(cfg, data_array, label_array, num_round, param)`` (the reference defines both, the
    second shadowing the first; ``wrapper/cxxnet.py:281-312``)."""
    if isinstance(data, DataIter):
        num_round, param = args[0], args[1]
        eval_data = args[2] if len(args) > 2 else kw.get("eval_data")
        net = Net(dev=kw.get("dev", ""), cfg=cfg)
        for k, v in (param.items() if isinstance(param, dict) else param):
            net.set_param(k, v)
        net.init_model()
        for r in range(num_round):
            net.start_round(r)
            data.before_first()
            scounter = 0
            while data.next():
                net.update(data)
                scounter += 1
                if scounter % 100 == 0:
                    print("[%d] %d batch passed" % (r, scounter))
            if eval_data is not None:
                sys.stderr.write(net.evaluate(eval_data, "eval") + "\n")
        return net
    label, num_round, param = args[0], args[1], args[2]
    net = Net(dev=kw.get("dev", ""), cfg=cfg)
    for k, v in (param.items() if isinstance(param, dict) else param):
        net.set_param(k, v)
    net.init_model()
    for r in range(num_round):
        print("Training in round %d" % r)
        net.start_round(r)
        net.update(data=data, label=label)
    return net
