"""Layer / node contracts shared by every layer kind.

Parity notes (reference ``src/layer/layer.h``): a node is a 4-D ``(batch, channel, height,
width)`` tensor, vectors are ``(batch,1,1,len)``; in the *reference backend* the same node buffer
holds the activation in forward and the gradient in backward, weight gradients accumulate
(``+=``) and are zeroed by the updater.  ``LayerParam`` mirrors ``src/layer/param.h:15-139``
including its raw 328-byte checkpoint image.

Two execution backends exist:

* ``ref``  - plain PyTorch fp32 NCHW restatement of the reference math (CPU or CUDA); this
  is the correctness oracle and the ``dev=cpu`` path.
* ``b200`` - NHWC bf16 activations, hand-written sm_100a kernels (``cxxnet_b200.ops``).
  Layers without a native implementation run their ``ref`` math on the device through a
  layout bridge (see ``executor.py``).
"""
from __future__ import annotations

import math
import struct
from typing import BinaryIO, Dict, List, Optional, Tuple

import numpy as np
import torch

from .. import model_io as mio


class Node:
    """One activation node of the graph."""

    __slots__ = ("shape", "data", "name", "index", "act", "grad", "f32", "alias_of", "loss_input", "wpitch", "tp_rows", "s2d")

    def __init__(self, name: str = "", index: int = 0):
        self.shape: List[int] = [0, 0, 0, 0]  # (b, c, h, w)
        self.data: Optional[torch.Tensor] = None  # ref backend buffer (activation, then gradient)
        self.name = name
        self.index = index
        # b200 backend buffers
        self.act = None    # ops.DevMat (bf16)
        self.grad = None   # ops.DevMat (bf16)
        self.f32 = None    # fp32 [b, ld] buffer for loss nodes
        self.alias_of = None
        self.loss_input = False
        self.wpitch = 0    # b200: pixels per stored image row when it differs from shape[3] (0 = dense)
        self.s2d = 0       # b200: 4 = stored space-to-depth, [N][ceil(H/4)][ceil(W/4)][64] (cxn_nchw_to_s2d4)
        # b200, several ranks: > 0 when the node lives between two column-sharded fully connected layers and is stored
        # as [global batch = tp_rows][this rank's columns] instead of [batch slice][all columns] (parallel/tensor_parallel.py)
        self.tp_rows = 0

    @property
    def is_mat(self) -> bool:
        return self.shape[1] == 1 and self.shape[2] == 1

    def mat(self) -> torch.Tensor:
        assert self.is_mat, "node is not a matrix node (b,1,1,n)"
        return self.data.view(self.shape[0], self.shape[3])


class LabelInfo:
    """Per-batch label fields (reference ``layer.h`` LabelInfo / LabelRecord)."""

    def __init__(self, name2findex: Dict[str, int]):
        self.name2findex = name2findex
        self.fields: List[torch.Tensor] = []  # each [b, width] float32


_LP_FMT = "<ififfiiiiiiiiiiiii64i"
assert struct.calcsize(_LP_FMT) == 328


class LayerParam:
    def __init__(self):
        self.num_hidden = 0
        self.init_sigma = 0.01
        self.init_sparse = 10
        self.init_uniform = -1.0
        self.init_bias = 0.0
        self.num_channel = 0
        self.random_type = 0
        self.num_group = 1
        self.kernel_height = 0
        self.kernel_width = 0
        self.stride = 1
        self.pad_y = 0
        self.pad_x = 0
        self.no_bias = 0
        self.temp_col_max = 64 << 18
        self.silent = 0
        self.num_input_channel = 0
        self.num_input_node = 0

    def set_param(self, name: str, val: str) -> None:
        if name == "init_sigma": self.init_sigma = float(val)
        if name == "init_uniform": self.init_uniform = float(val)
        if name == "init_bias": self.init_bias = float(val)
        if name == "init_sparse": self.init_sparse = int(val)
        if name == "random_type":
            if val == "gaussian": self.random_type = 0
            elif val in ("uniform", "xavier"): self.random_type = 1
            elif val == "kaiming": self.random_type = 2
            else: raise ValueError("invalid random_type %s" % val)
        if name == "nhidden": self.num_hidden = int(val)
        if name == "nchannel": self.num_channel = int(val)
        if name == "ngroup": self.num_group = int(val)
        if name == "kernel_size": self.kernel_width = self.kernel_height = int(val)
        if name == "kernel_height": self.kernel_height = int(val)
        if name == "kernel_width": self.kernel_width = int(val)
        if name == "stride": self.stride = int(val)
        if name == "pad": self.pad_y = self.pad_x = int(val)
        if name == "pad_y": self.pad_y = int(val)
        if name == "pad_x": self.pad_x = int(val)
        if name == "no_bias": self.no_bias = int(val)
        if name == "silent": self.silent = int(val)
        if name == "temp_col_max": self.temp_col_max = int(val) << 18

    def pack(self) -> bytes:
        return struct.pack(_LP_FMT, self.num_hidden, self.init_sigma, self.init_sparse, self.init_uniform,
                           self.init_bias, self.num_channel, self.random_type, self.num_group, self.kernel_height,
                           self.kernel_width, self.stride, self.pad_y, self.pad_x, self.no_bias, self.temp_col_max,
                           self.silent, self.num_input_channel, self.num_input_node, *([0] * 64))

    def unpack(self, b: bytes) -> None:
        v = struct.unpack(_LP_FMT, b)
        (self.num_hidden, self.init_sigma, self.init_sparse, self.init_uniform, self.init_bias, self.num_channel,
         self.random_type, self.num_group, self.kernel_height, self.kernel_width, self.stride, self.pad_y, self.pad_x,
         self.no_bias, self.temp_col_max, self.silent, self.num_input_channel, self.num_input_node) = v[:18]

    def save(self, fo: BinaryIO) -> None:
        fo.write(self.pack())

    def load(self, fi: BinaryIO) -> None:
        b = fi.read(328)
        if len(b) != 328:
            raise mio.ModelFormatError("invalid model file: LayerParam")
        self.unpack(b)

    def rand_init_weight(self, gen: torch.Generator, shape, in_num: int, out_num: int) -> torch.Tensor:
        """reference ``param.h:113-138``."""
        if self.random_type == 0:
            return torch.randn(shape, generator=gen) * self.init_sigma
        if self.random_type == 1:
            a = math.sqrt(3.0 / (in_num + out_num))
            if self.init_uniform > 0:
                a = self.init_uniform
            return (torch.rand(shape, generator=gen) * 2.0 - 1.0) * a
        if self.random_type == 2:
            if self.num_hidden > 0:
                sigma = math.sqrt(2.0 / self.num_hidden)
            else:
                sigma = math.sqrt(2.0 / (self.num_channel * self.kernel_width * self.kernel_height))
            return torch.randn(shape, generator=gen) * sigma
        raise ValueError("unsupported random_type")


class Layer:
    """Base class.  ``forward`` / ``backprop`` implement the *reference backend*."""

    type_name = "layer"
    # tag -> (weight attribute, gradient attribute); used to re-point a layer at arena storage
    weight_attrs = {"wmat": ("wmat", "gwmat"), "bias": ("bias", "gbias")}
    self_loop = False        # in == out node (dropout, loss, bias)
    allow_sharing = True
    has_b200 = False         # native sm_100a implementation available

    def __init__(self):
        self.param = LayerParam()
        self.rng: Optional[torch.Generator] = None
        self.label_info: Optional[LabelInfo] = None
        self.device = torch.device("cpu")
        self.layer_index = -1
        self.nodes_in: List[Node] = []
        self.nodes_out: List[Node] = []
        self.state: Dict[str, torch.Tensor] = {}

    # ---- configuration -------------------------------------------------
    def set_param(self, name: str, val: str) -> None:
        self.param.set_param(name, val)

    def init_connection(self, nodes_in: List[Node], nodes_out: List[Node]) -> None:
        raise NotImplementedError

    def on_batch_size_changed(self, nodes_in: List[Node], nodes_out: List[Node]) -> None:
        pass

    # ---- model ----------------------------------------------------------
    def init_model(self) -> None:
        pass

    def save_model(self, fo: BinaryIO) -> None:
        pass

    def load_model(self, fi: BinaryIO) -> None:
        pass

    def weights(self) -> List[Tuple[str, torch.Tensor, torch.Tensor]]:
        """(tag, weight, grad) triples as visited by the reference's IVisitor (reference layout)."""
        return []

    def to_device(self, device: torch.device) -> None:
        self.device = device
        for tag, w, g in self.weights():
            wa, ga = self.weight_attrs[tag]
            setattr(self, wa, w.to(device))
            setattr(self, ga, g.to(device))

    def rebind(self, tag: str, w: torch.Tensor, g: torch.Tensor) -> None:
        """Replace the storage of one weight (and its gradient) by the given tensors."""
        wa, ga = self.weight_attrs[tag]
        setattr(self, wa, w)
        setattr(self, ga, g)

    # ---- reference backend ---------------------------------------------
    def forward(self, is_train: bool, nodes_in: List[Node], nodes_out: List[Node]) -> None:
        raise NotImplementedError

    def backprop(self, prop_grad: bool, nodes_in: List[Node], nodes_out: List[Node]) -> None:
        raise NotImplementedError

    # ---- b200 backend ----------------------------------------------------
    def b200_setup(self, ex) -> None:
        pass

    def b200_forward(self, ex, is_train: bool) -> None:
        raise NotImplementedError

    def b200_backprop(self, ex, prop_grad: bool) -> None:
        raise NotImplementedError

    # weight layout bridge: reference layout <-> device-internal layout (default: identical)
    def weight_to_internal(self, tag: str, w: torch.Tensor) -> torch.Tensor:
        return w

    def weight_from_internal(self, tag: str, w: torch.Tensor) -> torch.Tensor:
        return w

    # ---- helpers ---------------------------------------------------------
    @staticmethod
    def _check_11(nodes_in, nodes_out, who: str) -> None:
        if len(nodes_in) != 1 or len(nodes_out) != 1:
            raise ValueError("%s: only support 1-1 connection" % who)

    def _save_tensor(self, fo: BinaryIO, t: torch.Tensor) -> None:
        mio.write_tensor(fo, t.detach().cpu().numpy())

    def _load_tensor(self, fi: BinaryIO, dim: int) -> torch.Tensor:
        return torch.from_numpy(mio.read_tensor(fi, dim)).to(self.device)


def pool_out_size(h: int, k: int, s: int) -> int:
    """ceil-style pooled size with border clipping (reference ``pooling_layer-inl.hpp:103-106``)."""
    return min(h - k + s - 1, h - 1) // s + 1
