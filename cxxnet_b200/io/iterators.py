"""Iterator implementations.

Instance level : ``MNISTIterator`` (batch level in the reference too), ``ImageIterator`` (img),
``ImageBinIterator`` (imgbin / imgbinx over BinaryPage files), ``SyntheticInstIterator``,
``AugmentIterator`` (+ ``ImageAugmenter`` affine warp).
Batch level    : ``BatchAdaptIterator``, ``ThreadBufferIterator``, ``DenseBufferIterator``,
``AttachTxtIterator``.
"""
from __future__ import annotations

import gzip
import math
import os
import queue
import struct
import threading
from typing import List, Optional

import numpy as np
import torch

from .. import model_io as mio
from .data import DataBatch, DataInst, IIterator

PAGE_INTS = 64 << 18  # BinaryPage: 64 Mi ints... 64<<18 ints = 64 MiB (reference src/utils/io.h:257)


def _decode_image(buf: bytes, native_first: bool = True) -> np.ndarray:
    """bytes -> float32 RGB planar (3,h,w), as the reference's imdecode + BGR->RGB split.  JPEGs go through the
    native decoder (csrc/runtime/jpeg_decoder.cpp, the counterpart of the reference's utils/decoder.h); other
    formats, and JPEG variants it does not cover, through OpenCV or Pillow.  Raw records (`im2bin ... raw=1`,
    csrc/runtime/raw_image.cpp) are unpacked directly."""
    if buf[:4] == b"CXNR" and len(buf) >= 12:
        w, h = int.from_bytes(buf[4:6], "little"), int.from_bytes(buf[6:8], "little")
        px = np.frombuffer(buf, dtype=np.uint8, count=w * h * 3, offset=12).reshape(h, w, 3)
        return np.ascontiguousarray(px.transpose(2, 0, 1), dtype=np.float32)
    if native_first and os.environ.get("CXN_NATIVE_JPEG", "1") != "0":
        from . import native
        arr = native.image_decode(buf)
        if arr is not None:
            return arr
    try:
        import cv2
        arr = cv2.imdecode(np.frombuffer(buf, dtype=np.uint8), 1)
        if arr is None:
            raise ValueError("decoding fail")
        return np.ascontiguousarray(arr[:, :, ::-1].transpose(2, 0, 1)).astype(np.float32)
    except ImportError:
        from PIL import Image
        import io as _io
        im = Image.open(_io.BytesIO(buf)).convert("RGB")
        return np.asarray(im, dtype=np.float32).transpose(2, 0, 1).copy()


class RandomSampler:
    """uniform sampler + Fisher-Yates shuffle (reference ``src/utils/random.h:16-52``)."""

    def __init__(self, seed: int = 0):
        self.rs = np.random.RandomState(seed & 0x7FFFFFFF)

    def seed(self, s: int):
        self.rs = np.random.RandomState(s & 0x7FFFFFFF)

    def next_double(self) -> float:
        return float(self.rs.random_sample())

    def next_uint32(self, n: int) -> int:
        return int(self.next_double() * n) if n > 0 else 0

    def shuffle(self, arr):
        for i in range(len(arr) - 1, 0, -1):
            j = self.next_uint32(i + 1)
            arr[i], arr[j] = arr[j], arr[i]


# ----------------------------------------------------------------------------- mnist
class MNISTIterator(IIterator):
    """idx(.gz) reader, ``/256`` normalisation, ``input_flat``, whole-set shuffle at init, last
    partial batch dropped (reference ``src/io/iter_mnist-inl.hpp:14-156``).  With ``synthetic = 1``
    (or missing files and ``synthetic = auto``) a deterministic 10-class surrogate of the same
    shapes is generated instead."""

    def __init__(self):
        self.silent = 0
        self.mode = 1
        self.shuffle = 0
        self.inst_offset = 0
        self.path_img = self.path_label = ""
        self.batch_size = 0
        self.rnd = RandomSampler(0)
        self.synthetic = "auto"
        self.num_synthetic = 6000
        self.out = DataBatch()
        self.loc = 0

    def set_param(self, name, val):
        if name == "silent": self.silent = int(val)
        if name == "batch_size": self.batch_size = int(val)
        if name == "input_flat": self.mode = int(val)
        if name == "shuffle": self.shuffle = int(val)
        if name == "index_offset": self.inst_offset = int(val)
        if name == "path_img": self.path_img = val
        if name == "path_label": self.path_label = val
        if name == "seed_data": self.rnd.seed(int(val))
        if name == "synthetic": self.synthetic = val
        if name == "num_synthetic": self.num_synthetic = int(val)

    @staticmethod
    def _open(path):
        return gzip.open(path, "rb") if path.endswith(".gz") else open(path, "rb")

    def _load(self):
        with self._open(self.path_img) as f:
            _, n, r, c = struct.unpack(">iiii", f.read(16))
            img = np.frombuffer(f.read(n * r * c), dtype=np.uint8).reshape(n, r, c).astype(np.float32)
        with self._open(self.path_label) as f:
            _, n2 = struct.unpack(">ii", f.read(8))
            lab = np.frombuffer(f.read(n2), dtype=np.uint8).astype(np.float32)
        return img * (1.0 / 256.0), lab

    def _synth(self):
        """10 fixed class prototypes + noise, 28x28, values in [0,1): linearly separable enough for
        the MLP / LeNet configs to reach low error, deterministic for a given seed."""
        rs = np.random.RandomState(1234)
        protos = (rs.rand(10, 28, 28) < 0.25).astype(np.float32)
        rs2 = np.random.RandomState(4321 + self.inst_offset)
        lab = rs2.randint(0, 10, size=self.num_synthetic).astype(np.float32)
        img = protos[lab.astype(int)] * 0.8 + rs2.rand(self.num_synthetic, 28, 28).astype(np.float32) * 0.3
        return np.clip(img, 0, 255.0 / 256.0).astype(np.float32), lab

    def init(self):
        have = os.path.exists(self.path_img) and os.path.exists(self.path_label)
        if self.synthetic == "1" or (self.synthetic == "auto" and not have):
            self.img, self.labels = self._synth()
            src = "synthetic"
        else:
            self.img, self.labels = self._load()
            src = self.path_img
        n = self.img.shape[0]
        self.inst = np.arange(n, dtype=np.uint32) + self.inst_offset
        if self.shuffle:
            order = list(range(n))
            self.rnd.shuffle(order)
            order = np.asarray(order)
            self.img, self.labels, self.inst = self.img[order], self.labels[order], self.inst[order]
        r, c = self.img.shape[1], self.img.shape[2]
        self.shape = (self.batch_size, 1, 1, r * c) if self.mode == 1 else (self.batch_size, 1, r, c)
        self.data_t = torch.from_numpy(np.ascontiguousarray(self.img))
        self.label_t = torch.from_numpy(np.ascontiguousarray(self.labels)).view(-1, 1)
        if torch.cuda.is_available():
            self.data_t, self.label_t = self.data_t.pin_memory(), self.label_t.pin_memory()
        self.out.batch_size = self.batch_size
        if self.silent == 0:
            print("MNISTIterator: load %u images (%s), shuffle=%d, shape=%u,%u,%u,%u" % ((n, src, self.shuffle) + self.shape))

    def before_first(self):
        self.loc = 0

    def next(self):
        if self.loc + self.batch_size <= self.img.shape[0]:
            a, b = self.loc, self.loc + self.batch_size
            self.out.data = self.data_t[a:b].view(self.shape)
            self.out.label = self.label_t[a:b]
            self.out.inst_index = self.inst[a:b]
            self.out.num_batch_padd = 0
            self.loc = b
            return True
        return False

    def value(self):
        return self.out


# ----------------------------------------------------------------------------- instance sources
class SyntheticInstIterator(IIterator):
    """Deterministic learnable instances of any shape (stands in for img / imgbin sources).

    keys: ``input_shape``, ``num_inst`` (1024), ``num_class`` (10), ``label_width`` (1),
    ``source_shape`` (c,h,w of the raw image before cropping; default = input_shape),
    ``seed_data``, ``value_scale`` (255 -> image-like range).
    """

    def __init__(self):
        self.shape = (1, 1, 1)
        self.src_shape = None
        self.num_inst = 1024
        self.num_class = 10
        self.label_width = 1
        self.seed = 0
        self.value_scale = 1.0
        self.i = 0
        self.out = DataInst()

    def set_param(self, name, val):
        if name == "input_shape": self.shape = tuple(int(x) for x in val.split(","))
        if name == "source_shape": self.src_shape = tuple(int(x) for x in val.split(","))
        if name == "num_inst": self.num_inst = int(val)
        if name == "num_class": self.num_class = int(val)
        if name == "label_width": self.label_width = int(val)
        if name == "seed_data": self.seed = int(val)
        if name == "value_scale": self.value_scale = float(val)

    def init(self):
        shp = self.src_shape or self.shape
        rs = np.random.RandomState(777 + self.seed)
        # low-rank class prototypes keep memory small for large images
        self.proto_a = rs.randn(self.num_class, shp[0], shp[1], 1).astype(np.float32)
        self.proto_b = rs.randn(self.num_class, shp[0], 1, shp[2]).astype(np.float32)
        self.shp = shp

    def before_first(self):
        self.i = 0

    def next(self):
        if self.i >= self.num_inst:
            return False
        rs = np.random.RandomState((self.seed * 1000003 + self.i) & 0x7FFFFFFF)
        k = rs.randint(0, self.num_class)
        x = self.proto_a[k] * self.proto_b[k] + 0.5 * rs.randn(*self.shp).astype(np.float32)
        self.out.data = (x * self.value_scale).astype(np.float32)
        lab = np.zeros(self.label_width, dtype=np.float32)
        lab[0] = k
        self.out.label = lab
        self.out.index = self.i
        self.i += 1
        return True

    def value(self):
        return self.out


class ImageIterator(IIterator):
    """image list ``index \\t label.. \\t path`` + decode (reference ``src/io/iter_img-inl.hpp:16-135``)."""

    def __init__(self):
        self.silent = 0
        self.shuffle = 0
        self.label_width = 1
        self.path_imglst = self.path_imgdir = ""
        self.raw_mode = False    # next_raw() hands over the file's bytes; the batch stage decodes (native threads)
        self.out = DataInst()

    def set_param(self, name, val):
        if name == "image_list": self.path_imglst = val
        if name == "image_root": self.path_imgdir = val
        if name == "silent": self.silent = int(val)
        if name == "shuffle": self.shuffle = int(val)
        if name == "label_width": self.label_width = int(val)

    def init(self):
        self.index_list, self.labels, self.filenames = [], [], []
        with open(self.path_imglst) as f:
            for line in f:
                t = line.split()
                if len(t) < 2 + self.label_width:
                    continue
                self.index_list.append(int(t[0]))
                self.labels.append([float(x) for x in t[1:1 + self.label_width]])
                self.filenames.append(t[1 + self.label_width])
        if self.silent == 0:
            print("ImageIterator:image_list=%s" % self.path_imglst)
        self.order = list(range(len(self.index_list)))
        self.rs = np.random.RandomState(0)
        self.before_first()

    def before_first(self):
        self.data_index = 0
        if self.shuffle:
            self.rs.shuffle(self.order)

    def next_raw(self):
        """(index, label, encoded bytes) of the next image, or None at the end (see ImageBinIterator.next_raw)"""
        if self.data_index >= len(self.order):
            return None
        i = self.order[self.data_index]
        self.data_index += 1
        with open(self.path_imgdir + self.filenames[i], "rb") as f:
            return self.index_list[i], np.asarray(self.labels[i], dtype=np.float32), f.read()

    def next(self):
        if self.data_index >= len(self.order):
            return False
        i = self.order[self.data_index]
        with open(self.path_imgdir + self.filenames[i], "rb") as f:
            self.out.data = _decode_image(f.read())
        self.out.index = self.index_list[i]
        self.out.label = np.asarray(self.labels[i], dtype=np.float32)
        self.data_index += 1
        return True

    def value(self):
        return self.out


class BinaryPage:
    """64 MiB page of packed blobs (reference ``src/utils/io.h:254-326``): ``data[0]`` = count,
    ``data[r+2]`` = end offset of blob r measured back from the end of the page."""

    def __init__(self, buf: Optional[bytes] = None):
        self.nbytes = PAGE_INTS * 4
        if buf is None:
            self.raw = bytearray(self.nbytes)
        else:
            assert len(buf) == self.nbytes
            self.raw = buf
        self.head = np.frombuffer(self.raw, dtype="<i4")

    def size(self) -> int:
        return int(self.head[0])

    def get(self, r: int) -> bytes:
        end, beg = int(self.head[r + 2]), int(self.head[r + 1])
        return bytes(self.raw[self.nbytes - end:self.nbytes - beg])

    def view(self, r: int) -> memoryview:
        """blob r without a copy: a view that keeps the page's buffer alive for as long as it is referenced"""
        end, beg = int(self.head[r + 2]), int(self.head[r + 1])
        return memoryview(self.raw)[self.nbytes - end:self.nbytes - beg]

    def free_bytes(self) -> int:
        n = self.size()
        return (PAGE_INTS - (n + 2)) * 4 - int(self.head[n + 1])

    def push(self, blob: bytes) -> bool:
        if self.free_bytes() < len(blob) + 4:
            return False
        n = self.size()
        head = np.frombuffer(self.raw, dtype="<i4")
        end = int(head[n + 1]) + len(blob)
        head[n + 2] = end
        self.raw[self.nbytes - end:self.nbytes - end + len(blob)] = blob
        head[0] = n + 1
        return True

    def clear(self):
        self.raw[:] = bytes(self.nbytes)


class ImageBinIterator(IIterator):
    """``imgbin`` / ``imgbinx``: packed BinaryPage file(s) + list file(s), background page
    prefetch, multi-file ``image_conf_prefix`` / ``image_conf_ids`` and worker sharding by
    ``dist_num_worker`` / ``dist_worker_rank`` / env ``PS_RANK`` (reference
    ``src/io/iter_thread_imbin-inl.hpp:16-283``, ``iter_thread_imbin_x-inl.hpp:18-395``).
    ``two_stage`` adds the decode-ahead thread and in-page shuffling of imgbinx.  When the native
    reader (``libcxxnet_io.so``) is built it does the page IO; otherwise a Python thread does.
    """

    def __init__(self, two_stage: bool):
        self.two_stage = two_stage
        self.raw_mode = False    # see next_raw
        self.silent = 0
        self.label_width = 1
        self.shuffle = 0
        self.dist_num_worker = 0
        self.dist_worker_rank = 0
        self.path_imgbin: List[str] = []
        self.path_imglst: List[str] = []
        self.img_conf_prefix = self.img_conf_ids = ""
        self.out = DataInst()
        self.rs = np.random.RandomState(0)
        self._q = None
        self._thr = None
        self.use_native = 1

    def set_param(self, name, val):
        if name == "native_reader": self.use_native = int(val)
        if name == "image_list": self.path_imglst.append(val)
        if name == "image_bin": self.path_imgbin.append(val)
        if name == "image_conf_prefix": self.img_conf_prefix = val
        if name == "image_conf_ids": self.img_conf_ids = val
        if name == "dist_num_worker": self.dist_num_worker = int(val)
        if name == "dist_worker_rank": self.dist_worker_rank = int(val)
        if name == "silent": self.silent = int(val)
        if name == "shuffle": self.shuffle = int(val)
        if name == "seed_data": self.rs = np.random.RandomState(int(val))
        if name == "label_width": self.label_width = int(val)

    def _parse_image_conf(self):
        r = os.environ.get("PS_RANK")
        if r is not None:
            self.dist_worker_rank = int(r)
        if not self.img_conf_prefix:
            return
        assert not self.path_imglst and not self.path_imgbin, \
            "you can either set image_conf_prefix or image_bin/image_list"
        lb, ub = [int(x) for x in self.img_conf_ids.split("-")]
        n = ub + 1 - lb
        if self.dist_num_worker > 1:
            step = (n + self.dist_num_worker - 1) // self.dist_num_worker
            begin = min(self.dist_worker_rank * step, n) + lb
            end = min((self.dist_worker_rank + 1) * step, n) + lb
            lb, ub = begin, end - 1
            assert lb <= ub, "ThreadImagePageIterator: too many workers such that idlist cannot be divided between them"
        for i in range(lb, ub + 1):
            p = self.img_conf_prefix % i
            self.path_imglst.append(p + ".lst")
            self.path_imgbin.append(p + ".bin")

    def init(self):
        self._parse_image_conf()
        assert len(self.path_imgbin) == len(self.path_imglst), "List/Bin number not consist"
        if self.silent == 0:
            print("ThreadImagePageIterator:image_list=%s, bin=%s" % (",".join(self.path_imglst), ",".join(self.path_imgbin)))
        self.before_first()

    def _records_native(self):
        """file-order records with the pages streamed by the native reader thread."""
        from . import native
        rd = native.PageReader(list(self.path_imgbin), 4)
        page, r = None, 0
        try:
            for fi in range(len(self.path_imglst)):
                with open(self.path_imglst[fi]) as fl:
                    for ln in fl:
                        t = ln.split()
                        if not t:
                            continue
                        while page is None or r >= page.size():
                            # a whole page into a buffer of its own; records are views of it (no per-record copy,
                            # the buffer lives as long as any of its records is referenced)
                            buf = np.empty(PAGE_INTS * 4, dtype=np.uint8)
                            if not rd.next_page(buf):
                                return
                            page, r = BinaryPage(buf), 0
                        blob = page.view(r)
                        r += 1
                        yield (int(t[0]), np.asarray([float(x) for x in t[1:1 + self.label_width]], dtype=np.float32),
                               blob)
        finally:
            rd.close()

    def _records(self):
        """generator of (index, label, blob) in file order, pages read ahead by a worker thread."""
        from . import native
        if not (self.two_stage and self.shuffle) and native.lib() is not None and self.use_native:
            yield from self._records_native()
            return
        order = list(range(len(self.path_imgbin)))
        if self.two_stage and self.shuffle:
            self.rs.shuffle(order)
        for fi in order:
            with open(self.path_imglst[fi]) as fl:
                lines = [ln.split() for ln in fl if ln.strip()]
            li = 0
            with open(self.path_imgbin[fi], "rb") as fb:
                while True:
                    buf = fb.read(PAGE_INTS * 4)
                    if len(buf) < PAGE_INTS * 4:
                        break
                    page = BinaryPage(buf)
                    ids = list(range(page.size()))
                    recs = []
                    for r in ids:
                        if li >= len(lines):
                            break
                        t = lines[li]
                        li += 1
                        recs.append((int(t[0]), np.asarray([float(x) for x in t[1:1 + self.label_width]],
                                                           dtype=np.float32), page.view(r)))
                    if self.two_stage and self.shuffle:
                        self.rs.shuffle(recs)
                    for rec in recs:
                        yield rec

    def _producer(self, q, stop):
        try:
            if self.two_stage and not self.raw_mode:
                # imgbinx decode stage: groups of blobs are decoded by native worker threads (GIL released)
                from . import native
                group = []

                def flush():
                    imgs = native.jpeg_decode_many([r[2] for r in group]) if native.lib() is not None \
                        else [None] * len(group)
                    for r, im in zip(group, imgs):
                        q.put((r[0], r[1], im if im is not None else _decode_image(r[2], native_first=False)))
                    group.clear()
                for rec in self._records():
                    if stop.is_set():
                        return
                    group.append(rec)
                    if len(group) >= 32:
                        flush()
                flush()
                return
            for rec in self._records():
                if stop.is_set():
                    return
                q.put(rec)
        finally:
            q.put(None)

    def before_first(self):
        self.close()
        self._q = queue.Queue(maxsize=256 if self.two_stage else 1024)
        self._stop = threading.Event()
        self._thr = threading.Thread(target=self._producer, args=(self._q, self._stop), daemon=True)
        self._thr.start()

    def close(self):
        if self._thr is not None:
            self._stop.set()
            while self._thr.is_alive():
                try:
                    self._q.get(timeout=0.05)
                except queue.Empty:
                    pass
            self._thr = None

    def next(self):
        item = self._q.get()
        if item is None:
            self._q.put(None)
            return False
        idx, lab, payload = item
        self.out.index, self.out.label = idx, lab
        self.out.data = payload if (self.two_stage and not self.raw_mode) else _decode_image(payload)
        return True

    def next_raw(self):
        """(index, label, encoded bytes) of the next record, or None at the end; needs ``raw_mode`` (set
        before ``before_first``), in which the producer thread hands over blobs without decoding them - the
        batch stage then decodes a whole batch on native worker threads"""
        assert self.raw_mode, "next_raw needs raw_mode"
        item = self._q.get()
        if item is None:
            self._q.put(None)
            return None
        return item

    def value(self):
        return self.out


# ----------------------------------------------------------------------------- augmentation
class ImageAugmenter:
    """affine rotate / shear / scale / aspect warp + crop (reference ``src/io/image_augmenter-inl.hpp:13-204``).
    The warp is an inverse-mapped resample in numpy: bicubic by default, as the reference's OpenCV ``warpAffine`` call
    (``cv::INTER_CUBIC``, ``image_augmenter-inl.hpp:105-108``: the Keys kernel with a = -0.75, constant border =
    ``fill_value``), in float arithmetic where OpenCV uses 1/32-pixel fixed-point tables; ``warp_interp = bilinear``
    selects the cheaper two-by-two kernel."""

    def __init__(self):
        self.shape = (0, 0, 0)
        self.rand_crop = 0
        self.max_rotate_angle = 0
        self.max_aspect_ratio = 0.0
        self.max_shear_ratio = 0.0
        self.min_crop_size = self.max_crop_size = -1
        self.rotate = -1
        self.max_random_scale = self.min_random_scale = 1.0
        self.min_img_size, self.max_img_size = 0.0, 1e10
        self.fill_value = 255
        self.rotate_list: List[int] = []
        self.bicubic = True

    def set_param(self, name, val):
        if name == "input_shape": self.shape = tuple(int(x) for x in val.split(","))
        if name == "rand_crop": self.rand_crop = int(val)
        if name == "max_rotate_angle": self.max_rotate_angle = int(float(val))
        if name == "max_shear_ratio": self.max_shear_ratio = float(val)
        if name == "max_aspect_ratio": self.max_aspect_ratio = float(val)
        if name == "min_crop_size": self.min_crop_size = int(val)
        if name == "max_crop_size": self.max_crop_size = int(val)
        if name == "min_random_scale": self.min_random_scale = float(val)
        if name == "max_random_scale": self.max_random_scale = float(val)
        if name == "min_img_size": self.min_img_size = float(val)
        if name == "max_img_size": self.max_img_size = float(val)
        if name == "fill_value": self.fill_value = int(val)
        if name == "rotate": self.rotate = int(val)
        if name == "rotate_list": self.rotate_list = [int(x) for x in val.split(",") if x]
        if name == "warp_interp":
            if val not in ("bicubic", "bilinear"):
                raise ValueError("warp_interp must be bicubic or bilinear")
            self.bicubic = val == "bicubic"

    def need_process(self) -> bool:
        if self.max_rotate_angle > 0 or self.max_shear_ratio > 0.0 or self.rotate > 0 or self.rotate_list:
            return True
        return self.min_crop_size > 0 and self.max_crop_size > 0

    def draw(self, rows: int, cols: int, rnd: RandomSampler):
        """the random decisions of one warp for a rows x cols image, in the reference's order: returns
        (Ai, new_h, new_w, y, x, cropped) - the inverse affine map as a 3x3 float64 matrix, the size of the warped
        image, and the origin of the input_shape crop taken from it (cropped is False when the warped image is
        smaller than input_shape and is passed on whole)"""
        s = rnd.next_double() * self.max_shear_ratio * 2 - self.max_shear_ratio
        angle = rnd.next_uint32(self.max_rotate_angle * 2) - self.max_rotate_angle
        if self.rotate > 0:
            angle = self.rotate
        if self.rotate_list:
            angle = self.rotate_list[rnd.next_uint32(len(self.rotate_list) - 1)]
        a, b = math.cos(angle / 180.0 * math.pi), math.sin(angle / 180.0 * math.pi)
        scale = rnd.next_double() * (self.max_random_scale - self.min_random_scale) + self.min_random_scale
        ratio = rnd.next_double() * self.max_aspect_ratio * 2 - self.max_aspect_ratio + 1
        hs = 2 * scale / (1 + ratio)
        ws = ratio * hs
        new_w = int(max(self.min_img_size, min(self.max_img_size, scale * cols)))
        new_h = int(max(self.min_img_size, min(self.max_img_size, scale * rows)))
        M = np.array([[hs * a - s * b * ws, hs * b + s * a * ws, 0.0], [-b * ws, a * ws, 0.0]], dtype=np.float64)
        M[0, 2] = (new_w - (M[0, 0] * cols + M[0, 1] * rows)) / 2
        M[1, 2] = (new_h - (M[1, 0] * cols + M[1, 1] * rows)) / 2
        # dst(x,y) = src(M^-1 [x,y,1])
        Ai = np.linalg.inv(np.vstack([M, [0, 0, 1]]))
        y, x = new_h - self.shape[2], new_w - self.shape[1]
        if y < 0 or x < 0:
            return Ai, new_h, new_w, 0, 0, False
        if self.rand_crop != 0:
            y, x = rnd.next_uint32(y + 1), rnd.next_uint32(x + 1)
        else:
            y, x = y // 2, x // 2
        return Ai, new_h, new_w, y, x, True

    @staticmethod
    def cubic_weights(t):
        """the four Keys (a = -0.75) weights of taps -1, 0, 1, 2 for fraction t (float32, this operation order - the
        native stage csrc/runtime/batch_pipeline.cpp evaluates the same expressions)"""
        A = np.float32(-0.75)
        one = np.float32(1.0)
        t1 = t + one
        c0 = ((A * t1 - np.float32(5.0) * A) * t1 + np.float32(8.0) * A) * t1 - np.float32(4.0) * A
        c1 = ((A + np.float32(2.0)) * t - (A + np.float32(3.0))) * t * t + one
        u = one - t
        c2 = ((A + np.float32(2.0)) * u - (A + np.float32(3.0))) * u * u + one
        c3 = one - c0 - c1 - c2
        return c0, c1, c2, c3

    def warp(self, data: np.ndarray, params) -> np.ndarray:
        """inverse-mapped resample (bicubic or bilinear) of `data` (c, rows, cols) with the decisions of `draw`"""
        Ai, new_h, new_w, y, x, cropped = params
        c, rows, cols = data.shape
        ys, xs = np.mgrid[0:new_h, 0:new_w]
        sx = Ai[0, 0] * xs + Ai[0, 1] * ys + Ai[0, 2]
        sy = Ai[1, 0] * xs + Ai[1, 1] * ys + Ai[1, 2]
        x0, y0 = np.floor(sx).astype(int), np.floor(sy).astype(int)
        fx, fy = (sx - x0).astype(np.float32), (sy - y0).astype(np.float32)
        data = np.asarray(data, dtype=np.float32)
        fill = np.float32(self.fill_value)

        def tap(yy, xx):
            ok = (yy >= 0) & (yy < rows) & (xx >= 0) & (xx < cols)
            v = data[:, np.clip(yy, 0, rows - 1), np.clip(xx, 0, cols - 1)]
            return np.where(ok[None], v, fill)
        if self.bicubic:
            cx, cy = self.cubic_weights(fx), self.cubic_weights(fy)
            out = None
            for j in range(4):
                r = ((tap(y0 - 1 + j, x0 - 1) * cx[0] + tap(y0 - 1 + j, x0) * cx[1]) + tap(y0 - 1 + j, x0 + 1) * cx[2]) \
                    + tap(y0 - 1 + j, x0 + 2) * cx[3]
                out = r * cy[j] if out is None else out + r * cy[j]
            out = out.astype(np.float32)
        else:
            out = (tap(y0, x0) * (1 - fx) * (1 - fy) + tap(y0, x0 + 1) * fx * (1 - fy)
                   + tap(y0 + 1, x0) * (1 - fx) * fy + tap(y0 + 1, x0 + 1) * fx * fy).astype(np.float32)
        if not cropped:
            return out
        return out[:, y:y + self.shape[2], x:x + self.shape[1]]

    def process(self, data: np.ndarray, rnd: RandomSampler) -> np.ndarray:
        if not self.need_process():
            return data
        return self.warp(data, self.draw(data.shape[1], data.shape[2], rnd))


class AugmentIterator(IIterator):
    """mean image / mean value, crop, mirror, contrast / illumination jitter, scaling
    (reference ``src/io/iter_augment_proc-inl.hpp:21-246``)."""

    def __init__(self, base: IIterator):
        self.base = base
        self.shape = (0, 0, 0)
        self.rand_crop = self.rand_mirror = self.mirror = 0
        self.crop_y_start = self.crop_x_start = -1
        self.scale = 1.0
        self.silent = 0
        self.name_meanimg = ""
        self.mean_r = self.mean_g = self.mean_b = 0.0
        self.max_random_illumination = self.max_random_contrast = 0.0
        self.rnd = RandomSampler(0)
        self.aug = ImageAugmenter()
        self.meanimg = None
        self.meanfile_ready = False
        self.out = DataInst()

    def set_param(self, name, val):
        self.base.set_param(name, val)
        if name == "input_shape": self.shape = tuple(int(x) for x in val.split(","))
        if name == "seed_data": self.rnd.seed(int(val))
        if name == "rand_crop": self.rand_crop = int(val)
        if name == "silent": self.silent = int(val)
        if name == "divideby": self.scale = 1.0 / float(val)
        if name == "scale": self.scale = float(val)
        if name == "image_mean": self.name_meanimg = val
        if name == "crop_y_start": self.crop_y_start = int(val)
        if name == "crop_x_start": self.crop_x_start = int(val)
        if name == "rand_mirror": self.rand_mirror = int(val)
        if name == "mirror": self.mirror = int(val)
        if name == "max_random_contrast": self.max_random_contrast = float(val)
        if name == "max_random_illumination": self.max_random_illumination = float(val)
        if name == "mean_value":
            vals = [float(x) for x in val.split(",")]
            if len(vals) == 1:             # one number for every plane (single-channel nets)
                vals = vals * 3
            self.mean_b, self.mean_g, self.mean_r = vals
        self.aug.set_param(name, val)

    def init(self):
        self.base.init()
        self.meanfile_ready = False
        if self.name_meanimg:
            if os.path.exists(self.name_meanimg):
                if self.silent == 0:
                    print("loading mean image from %s" % self.name_meanimg)
                with open(self.name_meanimg, "rb") as f:
                    self.meanimg = mio.read_tensor(f, 3)
                self.meanfile_ready = True
            else:
                self._create_mean_img()

    def before_first(self):
        self.base.before_first()

    def close(self):
        self.base.close()

    # ---- the random decisions of one instance, drawn in the reference's order (iter_augment_proc-inl.hpp:98-162)
    def mean_mode(self) -> int:
        """0 none, 1 mean_value per channel, 2 / 3 mean image (decided per instance by its size: 2 when the
        image has the mean image's size - subtract, then crop - else 3: crop, then subtract)"""
        if self.mean_r > 0 or self.mean_g > 0 or self.mean_b > 0:
            return 1
        if not self.meanfile_ready or not self.name_meanimg:
            return 0
        return 2

    def draw_params(self, src_h: int, src_w: int):
        """(y0, x0, flip, contrast, illum) for a decoded image of the given size"""
        c, H, W = self.shape
        assert src_h >= H and src_w >= W, "Data size must be bigger than the input size to net."
        yy, xx = src_h - H, src_w - W
        if self.rand_crop != 0 and (yy != 0 or xx != 0):
            yy, xx = self.rnd.next_uint32(yy + 1), self.rnd.next_uint32(xx + 1)
        else:
            yy, xx = yy // 2, xx // 2
        if src_h != H and self.crop_y_start != -1: yy = self.crop_y_start
        if src_w != W and self.crop_x_start != -1: xx = self.crop_x_start
        contrast = self.rnd.next_double() * self.max_random_contrast * 2 - self.max_random_contrast + 1
        illum = self.rnd.next_double() * self.max_random_illumination * 2 - self.max_random_illumination
        flip = self.rand_mirror != 0 and self.rnd.next_double() < 0.5
        if self.mean_mode() != 0:
            flip = flip or self.mirror == 1
        return yy, xx, bool(flip), contrast, illum

    def apply_params(self, data: np.ndarray, params) -> np.ndarray:
        """NumPy statement of the per-instance arithmetic (also the oracle of the native batch path)"""
        yy, xx, flip, contrast, illum = params
        c, H, W = self.shape

        def crop(t):
            return t[:, yy:yy + H, xx:xx + W]
        mode = self.mean_mode()
        if mode == 1:
            data = data.copy()
            data[0] -= self.mean_b
            if data.shape[0] > 2:
                data[1] -= self.mean_g
                data[2] -= self.mean_r
            img = crop(data * contrast + illum)
            if flip:
                img = img[:, :, ::-1]
        elif mode == 0:
            img = crop(data)
            if flip:
                img = img[:, :, ::-1]
        elif data.shape == self.meanimg.shape:
            img = crop((data - self.meanimg) * contrast + illum)
            if flip:
                img = img[:, :, ::-1]
        else:
            img = crop(data)
            if flip:
                img = img[:, :, ::-1]
            img = (img - self.meanimg) * contrast + illum
        return np.ascontiguousarray(img * self.scale, dtype=np.float32)

    def _set_data(self, d: DataInst):
        self.out.label, self.out.index = d.label, d.index
        data = d.data
        if self.shape[0] == 1 and data.ndim == 3 and data.shape[0] == 3:
            # a single-channel net fed by an image source (decoders always deliver three planes; a grey-scale
            # file has them equal): the mean of the planes.  Not in the reference, whose image sources are
            # three-channel only.
            data = data.mean(axis=0, keepdims=True, dtype=np.float32)
        data = self.aug.process(data, self.rnd)
        c, H, W = self.shape
        if H == 1:
            self.out.data = (data * self.scale).astype(np.float32)
            return
        self.out.data = self.apply_params(data, self.draw_params(data.shape[1], data.shape[2]))

    def next(self):
        if not self.base.next():
            return False
        self._set_data(self.base.value())
        return True

    def value(self):
        return self.out

    def _create_mean_img(self):
        if self.silent == 0:
            print("cannot find %s: create mean image, this will take some time..." % self.name_meanimg)
        assert self.next(), "input iterator failed."
        mean = self.out.data.astype(np.float64).copy()
        cnt = 1
        while self.next():
            mean += self.out.data
            cnt += 1
        self.meanimg = (mean / cnt).astype(np.float32)
        d = os.path.dirname(self.name_meanimg)
        if d:
            os.makedirs(d, exist_ok=True)
        with open(self.name_meanimg, "wb") as f:
            mio.write_tensor(f, self.meanimg)
        if self.silent == 0:
            print("save mean image to %s.." % self.name_meanimg)
        self.meanfile_ready = True
        self.before_first()


# ----------------------------------------------------------------------------- batch level
class BatchAdaptIterator(IIterator):
    """DataInst -> DataBatch with ``round_batch`` wrap-around, ``num_batch_padd``, ``label_width``
    and ``test_skipread`` (reference ``src/io/iter_batch_proc-inl.hpp:16-133``)."""

    def __init__(self, base: IIterator):
        self.base = base
        self.batch_size = 0
        self.shape = (0, 0, 0)
        self.label_width = 1
        self.round_batch = 0
        self.num_overflow = 0
        self.silent = 0
        self.test_skipread = 0
        self.head = 1
        self.out = DataBatch()
        self.input_dtype = "fp32"     # bf16: batches leave the iterator as bf16 (half the H2D bytes; the b200
        self._half = None             # backend computes in bf16 anyway).  Not a reference option.
        self._u8 = None               # u8: raw pixels + DataBatch.norm, normalised where the batch enters the net
        self.native_pipeline = 1      # decode + augment + batch assembly on native threads where applicable
        self.decode_threads = 0       # 0 = min(8, cores)
        self._native = False
        self._pending = []            # native path: records of the batch being assembled
        self.slice_world, self.slice_rank = 1, 0

    def set_param(self, name, val):
        self.base.set_param(name, val)
        if name == "input_dtype":
            if val not in ("fp32", "bf16", "u8"):
                raise ValueError("input_dtype must be fp32, bf16 or u8")
            self.input_dtype = val
        if name == "batch_size": self.batch_size = int(val)
        if name == "input_shape": self.shape = tuple(int(x) for x in val.split(","))
        if name == "label_width": self.label_width = int(val)
        if name == "round_batch": self.round_batch = int(val)
        if name == "silent": self.silent = int(val)
        if name == "test_skipread": self.test_skipread = int(val)
        if name == "native_pipeline": self.native_pipeline = int(val)
        if name == "decode_threads": self.decode_threads = int(val)
        # one process per GPU: rank r of w only uses rows [r*ceil(B/w), (r+1)*ceil(B/w)) of every batch (trainer.py
        # _slice), so the native path decodes just those; the other rows of `data` are left as they are.  Set by the
        # task driver; labels, indices and the random decisions are still produced for every row.
        if name == "slice_world": self.slice_world = max(int(val), 1)
        if name == "slice_rank": self.slice_rank = int(val)

    def _native_chain(self) -> bool:
        """The fused native batch path (csrc/runtime/batch_pipeline.cpp) applies when the chain below is
        img / imgbin(x) -> augment on 3-channel images, and the library is built."""
        from . import native
        aug = self.base
        if not (self.native_pipeline and isinstance(aug, AugmentIterator)
                and isinstance(aug.base, (ImageBinIterator, ImageIterator))):
            return False
        if native.lib() is None or not hasattr(native.lib(), "cxnio_decode_augment_batch"):
            return False
        return not (len(self.shape) != 3 or self.shape[0] not in (1, 3) or self.shape[1] == 1)

    def init(self):
        chain = self._native_chain()
        if chain:
            # the source hands over encoded records from its first pass on (whoever reads it through next()
            # - the mean-image pass - gets them decoded on the spot), so no restart is needed later and a
            # shuffled source visits the data in the same order as without the native path
            self.base.base.raw_mode = True
        self.base.init()
        aug = self.base
        self._native = chain and not (aug.mean_mode() == 2 and (aug.meanimg is None
                                                                 or aug.meanimg.shape[0] != self.shape[0]))
        self.out.alloc_dense((self.batch_size,) + tuple(self.shape), self.batch_size, self.label_width, pin=True)
        self._data_np = self.out.data.numpy()
        self._label_np = self.out.label.numpy()
        self._u8 = None
        if self.input_dtype == "u8":
            # 8-bit batches: cropped / mirrored raw pixels cross PCIe (a quarter of the fp32 bytes), the net applies
            # (pixel - mean[c]) * scale on entry (DataBatch.norm).  That is only the same thing when the rest of the
            # augmentation is the identity on pixel values.
            ok = (self._native and self.shape[0] == 3 and aug.mean_mode() in (0, 1) and not aug.aug.need_process()
                  and aug.max_random_contrast == 0 and aug.max_random_illumination == 0)
            if not ok:
                raise ValueError("input_dtype = u8 needs the native batch path (img / imgbin source, 3 channels), "
                                 "mean_value or no mean, and no contrast / illumination / affine augmentation")
            self._u8 = DataBatch()
            self._u8.alloc_dense((self.batch_size,) + tuple(self.shape), self.batch_size, self.label_width, pin=True,
                                 dtype=torch.uint8)
            mean = [aug.mean_b, aug.mean_g, aug.mean_r] if aug.mean_mode() == 1 else [0.0, 0.0, 0.0]
            self._u8.norm = (mean, float(aug.scale))
            self._u8.label, self._u8.inst_index = self.out.label, self.out.inst_index
        if self.input_dtype == "bf16":
            self._half = DataBatch()
            self._half.alloc_dense((self.batch_size,) + tuple(self.shape), self.batch_size, self.label_width, pin=True,
                                   dtype=torch.bfloat16)
            self._half.label, self._half.inst_index = self.out.label, self.out.inst_index

    def before_first(self):
        if self.round_batch == 0 or self.num_overflow == 0:
            self.base.before_first()
        else:
            self.num_overflow = 0
        self.head = 1

    def close(self):
        self.base.close()

    def _put(self, top, d):
        self._label_np[top, :] = d.label[: self.label_width]
        self.out.inst_index[top] = d.index
        self._data_np[top] = d.data.reshape(self._data_np.shape[1:])

    # the two ways of taking one instance from below: decoded and augmented by the Python stages, or - native
    # path - as an encoded record that is only noted; _flush_native then processes the whole batch at once
    def _take(self, top) -> bool:
        if not self._native:
            if not self.base.next():
                return False
            self._put(top, self.base.value())
            return True
        rec = self.base.base.next_raw()
        if rec is None:
            return False
        self._label_np[top, :] = rec[1][: self.label_width]
        self.out.inst_index[top] = rec[0]
        self._pending.append((top, rec[2]))
        return True

    def _flush_native(self) -> None:
        """decode + augment every noted record into its batch slot (native worker threads)"""
        if not self._pending:
            return
        from . import native
        aug = self.base
        step = (self.batch_size + self.slice_world - 1) // self.slice_world
        lo, hi = min(self.slice_rank * step, self.batch_size), min((self.slice_rank + 1) * step, self.batch_size)
        items, tops = [], []
        for top, blob in self._pending:
            info = native.image_info(blob)
            pixels = None
            if info is None:                               # neither a JPEG the native decoder takes nor a raw record
                pixels = _decode_image(blob, native_first=False)
                if self.shape[0] == 1 and pixels.shape[0] == 3:
                    pixels = pixels.mean(axis=0, keepdims=True, dtype=np.float32)
                h, w = pixels.shape[1], pixels.shape[2]
            else:
                w, h = info[0], info[1]
            # drawn for every row: the stream of random decisions does not depend on how the batch is split over ranks
            affine = None
            if aug.aug.need_process():
                Ai, new_h, new_w, wy, wx, cropped = aug.aug.draw(h, w, aug.rnd)
                if not cropped:
                    raise ValueError("Data size must be bigger than the input size to net.")
                affine = (Ai, new_h, new_w, wy, wx, float(aug.aug.fill_value), 1 if aug.aug.bicubic else 0)
                params = tuple(aug.draw_params(self.shape[1], self.shape[2]))
            else:
                params = tuple(aug.draw_params(h, w))
            if lo <= top < hi:
                items.append((blob, pixels, h, w) + params + (affine,))
                tops.append(top)
        self._pending = []
        if not items:
            return
        mode = aug.mean_mode()
        mean = None
        if mode == 1:
            mean = np.asarray([aug.mean_b, aug.mean_g, aug.mean_r], dtype=np.float32)
        elif mode == 2:
            mean = aug.meanimg
        to_half = self._half is not None
        dst = self._half.data if to_half else self.out.data
        scale = aug.scale
        if self._u8 is not None:                           # raw pixels: the net normalises them (DataBatch.norm)
            dst, mode, mean, scale = self._u8.data, 0, None, 1.0
        first = tops[0]
        assert tops == list(range(first, first + len(items)))
        status = native.decode_augment_batch(items, self.shape, mode, mean, scale, dst[first:first + len(items)],
                                             to_half, self.decode_threads)
        for k, st in enumerate(status):
            if st != 0:                                    # e.g. a damaged file: the library decoders get a try
                blob, pixels, h, w = items[k][:4]
                data = pixels if pixels is not None else _decode_image(blob, native_first=False)
                if self.shape[0] == 1 and data.shape[0] == 3:
                    data = data.mean(axis=0, keepdims=True, dtype=np.float32)
                if self._u8 is not None:
                    y0, x0, flip = items[k][4:7]
                    img = data[:, y0:y0 + self.shape[1], x0:x0 + self.shape[2]]
                    img = img[:, :, ::-1] if flip else img
                    dst[first + k].copy_(torch.from_numpy(np.ascontiguousarray(np.clip(img + 0.5, 0, 255.5)).astype(np.uint8)))
                    continue
                if items[k][9] is not None:
                    Ai, new_h, new_w, wy, wx, _ = items[k][9]
                    data = aug.aug.warp(data, (Ai, new_h, new_w, wy, wx, True))
                res = torch.from_numpy(aug.apply_params(data, items[k][4:9]))
                dst[first + k].copy_(res)

    def next(self):
        self.out.num_batch_padd = 0
        if self.test_skipread != 0 and self.head == 0:
            return True
        self.head = 0
        if self.num_overflow != 0:
            return False
        top = 0
        while self._take(top):
            top += 1
            if top >= self.batch_size:
                self._flush_native()
                return True
        if top != 0:
            if self.round_batch != 0:
                self.num_overflow = 0
                self.base.before_first()
                while top < self.batch_size:
                    assert self._take(top), "number of input must be bigger than batch size"
                    top += 1
                    self.num_overflow += 1
                self.out.num_batch_padd = self.num_overflow
            else:
                self.out.num_batch_padd = self.batch_size - top
            self._flush_native()
            return True
        return False

    def value(self):
        assert self.head == 0, "must call Next to get value"
        if self._u8 is not None:
            self._u8.batch_size, self._u8.num_batch_padd = self.out.batch_size, self.out.num_batch_padd
            self._u8.extra_data = self.out.extra_data
            return self._u8
        if self._half is not None and self._native:
            # the native path wrote bf16 straight into the pinned half-precision batch
            self._half.batch_size, self._half.num_batch_padd = self.out.batch_size, self.out.num_batch_padd
            self._half.extra_data = self.out.extra_data
            return self._half
        if self._half is not None:
            self._half.data.copy_(self.out.data)      # fp32 -> bf16 on the producer side
            self._half.batch_size, self._half.num_batch_padd = self.out.batch_size, self.out.num_batch_padd
            self._half.extra_data = self.out.extra_data
            return self._half
        return self.out


class ThreadBufferIterator(IIterator):
    """double-buffered prefetch thread over a batch iterator (``buffer_size`` 2; reference
    ``src/io/iter_batch_proc-inl.hpp:136-224`` over ``utils/thread_buffer.h:22-202``).  Batches are
    copied into pinned slots so the trainer's H2D copy can be asynchronous."""

    def __init__(self, base: IIterator):
        self.base = base
        self.buffer_size = 2
        self.silent = 0
        self._thr = None
        self._cur = None

    def set_param(self, name, val):
        if name == "silent": self.silent = int(val)
        if name == "buffer_size": self.buffer_size = int(val)
        self.base.set_param(name, val)

    def init(self):
        self.base.init()
        if self.silent == 0:
            print("ThreadBufferIterator: buffer_size=%d" % self.buffer_size)
        self._slots = [DataBatch() for _ in range(self.buffer_size + 2)]
        self._start()

    def _start(self):
        self._free = queue.Queue()
        for s in self._slots:
            self._free.put(s)
        self._full = queue.Queue()
        self._stop = threading.Event()
        self._thr = threading.Thread(target=self._run, daemon=True)
        self._thr.start()

    def _run(self):
        self.base.before_first()
        while not self._stop.is_set():
            if not self.base.next():
                break
            slot = None
            while slot is None and not self._stop.is_set():
                try:
                    slot = self._free.get(timeout=0.05)
                except queue.Empty:
                    pass
            if slot is None:
                break
            slot.copy_from(self.base.value())
            self._full.put(slot)
        self._full.put(None)

    def _halt(self):
        if self._thr is not None:
            self._stop.set()
            while self._thr.is_alive():
                try:
                    self._full.get(timeout=0.05)
                except queue.Empty:
                    pass
            self._thr = None

    def before_first(self):
        # the producer already sits at the start of a fresh pass if nothing has been consumed
        if self._thr is not None and self._cur is None and not getattr(self, "_ended", False):
            return
        self._halt()
        self._cur = None
        self._ended = False
        self._start()

    def close(self):
        self._halt()
        self.base.close()

    def next(self):
        if getattr(self, "_ended", False):
            return False
        if self._cur is not None:
            self._free.put(self._cur)
            self._cur = None
        item = self._full.get()
        if item is None:
            self._ended = True
            return False
        self._cur = item
        return True

    def value(self):
        return self._cur


class DenseBufferIterator(IIterator):
    """cache the first ``max_nbatch`` batches in RAM and replay them (reference
    ``src/io/iter_mem_buffer-inl.hpp:16-75``)."""

    def __init__(self, base: IIterator):
        self.base = base
        self.max_nbatch = 100
        self.silent = 0
        self.buffer: List[DataBatch] = []
        self.data_index = 0

    def set_param(self, name, val):
        self.base.set_param(name, val)
        if name == "max_nbatch": self.max_nbatch = int(val)
        if name == "silent": self.silent = int(val)

    def init(self):
        self.base.init()
        self.base.before_first()
        while self.base.next():
            b = DataBatch()
            b.copy_from(self.base.value())
            self.buffer.append(b)
            if len(self.buffer) >= self.max_nbatch:
                break
        if self.silent == 0:
            print("DenseBufferIterator: load %d batches" % len(self.buffer))

    def before_first(self):
        self.data_index = 0

    def close(self):
        self.base.close()

    def next(self):
        if self.data_index < len(self.buffer):
            self.data_index += 1
            return True
        return False

    def value(self):
        return self.buffer[self.data_index - 1]


class AttachTxtIterator(IIterator):
    """join a text table (``dim`` header line, then ``id v1..vdim``) by ``inst_index`` into
    ``extra_data[0]`` (reference ``src/io/iter_attach_txt-inl.hpp:15-99``)."""

    def __init__(self, base: IIterator):
        self.base = base
        self.filename = ""
        self.batch_size = 0
        self.table = {}
        self.dim = 0
        self.out = DataBatch()

    def set_param(self, name, val):
        self.base.set_param(name, val)
        if name == "filename": self.filename = val
        if name == "batch_size": self.batch_size = int(val)

    def init(self):
        self.base.init()
        with open(self.filename) as f:
            toks = f.read().split()
        self.dim = int(toks[0])
        i = 1
        while i + self.dim < len(toks) + 0 and i < len(toks):
            key = int(toks[i])
            self.table[key] = np.asarray([float(x) for x in toks[i + 1:i + 1 + self.dim]], dtype=np.float32)
            i += 1 + self.dim

    def before_first(self):
        self.base.before_first()

    def close(self):
        self.base.close()

    def next(self):
        if not self.base.next():
            return False
        b = self.base.value()
        extra = torch.zeros((b.batch_size, 1, 1, self.dim), dtype=torch.float32)
        for r, idx in enumerate(b.inst_index[: b.batch_size]):
            v = self.table.get(int(idx))
            if v is not None:
                extra[r, 0, 0, :] = torch.from_numpy(v)
        self.out = b
        self.out.extra_data = [extra]
        return True

    def value(self):
        return self.out
