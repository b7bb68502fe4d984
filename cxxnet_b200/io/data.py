"""Data contracts and the iterator factory.

``DataInst`` / ``DataBatch`` / ``IIterator`` / ``create_iterator`` follow reference
``src/io/data.h:19-186`` and ``src/io/data.cpp:23-74``: an iterator *chain* is built in declaration
order from ``iter = <type>`` lines, every following ``name = value`` pair is offered to the
iterator built so far, and ``iter = end`` closes the block.

On top of the reference's kinds (mnist, img, imgbin, imgbinx, threadbuffer, membuffer,
attachtxt) there is ``iter = synthetic`` - a deterministic generator of learnable data of
any ``input_shape`` - and ``mnist`` / ``imgbin`` fall back to it when their files are absent
and ``synthetic = 1`` is set (there is no dataset access in the build environment).
"""
from __future__ import annotations

from typing import List, Optional

import numpy as np
import torch

from ..config import ConfigPairs


class DataInst:
    """one instance (reference ``data.h:57-70``)."""
    __slots__ = ("index", "label", "data")

    def __init__(self, index=0, label=None, data=None):
        self.index = index
        self.label = label      # np.float32 [label_width]
        self.data = data        # np.float32 [c, h, w]


class DataBatch:
    """a mini-batch (reference ``data.h:73-166``): ``data`` (b,c,h,w) fp32, ``label`` (b,label_width),
    ``inst_index`` (b,), ``num_batch_padd`` trailing padded rows, optional ``extra_data``."""

    def __init__(self):
        self.inst_index: Optional[np.ndarray] = None
        self.batch_size = 0
        self.num_batch_padd = 0
        self.label: Optional[torch.Tensor] = None
        self.data: Optional[torch.Tensor] = None
        self.extra_data: List[torch.Tensor] = []
        # 8-bit batches (`input_dtype = u8`): (per-channel means, scale) the net applies on entry,
        # x = (pixel - mean[c]) * scale; None for float batches, which arrive normalised
        self.norm = None

    def alloc_dense(self, shape, batch_size, label_width, pin=False, dtype=torch.float32):
        self.data = torch.zeros(tuple(shape), dtype=dtype)
        self.label = torch.zeros((batch_size, label_width), dtype=torch.float32)
        if pin and torch.cuda.is_available():
            self.data = self.data.pin_memory()
            self.label = self.label.pin_memory()
        self.inst_index = np.zeros(batch_size, dtype=np.uint32)
        self.batch_size = batch_size

    def copy_from(self, o: "DataBatch"):
        if self.data is None or self.data.shape != o.data.shape or self.data.dtype != o.data.dtype:
            self.alloc_dense(o.data.shape, o.batch_size, o.label.shape[1], pin=True, dtype=o.data.dtype)
        self.data.copy_(o.data)
        self.label.copy_(o.label)
        if o.inst_index is not None:
            self.inst_index[:] = o.inst_index
        self.batch_size = o.batch_size
        self.num_batch_padd = o.num_batch_padd
        self.norm = getattr(o, "norm", None)
        self.extra_data = [e.clone() for e in o.extra_data]


class IIterator:
    """reference ``data.h:19-54``."""

    def set_param(self, name: str, val: str) -> None:
        pass

    def init(self) -> None:
        pass

    def before_first(self) -> None:
        raise NotImplementedError

    def next(self) -> bool:
        raise NotImplementedError

    def value(self):
        raise NotImplementedError

    def close(self) -> None:
        pass

    def __iter__(self):
        self.before_first()
        while self.next():
            yield self.value()


def create_iterator(cfg: ConfigPairs) -> IIterator:
    from . import iterators as its
    it = None
    for name, val in cfg:
        if name == "iter":
            if val == "mnist":
                assert it is None, "mnist can not chain over other iterator"
                it = its.MNISTIterator()
                continue
            if val == "synthetic":
                assert it is None, "synthetic can not chain over other iterator"
                it = its.BatchAdaptIterator(its.AugmentIterator(its.SyntheticInstIterator()))
                continue
            if val == "imgbin":
                assert it is None, "image binary can not chain over other iterator"
                it = its.BatchAdaptIterator(its.AugmentIterator(its.ImageBinIterator(two_stage=False)))
                continue
            if val == "imgbinx":
                assert it is None, "image binary can not chain over other iterator"
                it = its.BatchAdaptIterator(its.AugmentIterator(its.ImageBinIterator(two_stage=True)))
                continue
            if val == "img":
                assert it is None, "image list iterator can not chain over other iterator"
                it = its.BatchAdaptIterator(its.AugmentIterator(its.ImageIterator()))
                continue
            if val == "threadbuffer":
                assert it is not None, "must specify input of threadbuffer"
                it = its.ThreadBufferIterator(it)
                continue
            if val == "membuffer":
                assert it is not None, "must specify input of memory buffer"
                it = its.DenseBufferIterator(it)
                continue
            if val == "attachtxt":
                assert it is not None, "must specify input of attach txt buffer"
                it = its.AttachTxtIterator(it)
                continue
            if val == "end":
                continue
            raise ValueError("unknown iterator type %s" % val)
        if it is not None:
            it.set_param(name, val)
    assert it is not None, "must specify iterator by iter=itername"
    return it
