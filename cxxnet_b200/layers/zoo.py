"""The layer zoo: reference-backend math for every layer kind of SURVEY.md section 2.2.

Each class restates, in plain PyTorch fp32 NCHW, what the corresponding reference layer
computes (file:line cited per class).  These are the oracle ("pairtest master") for the
sm_100a kernels and the ``dev=cpu`` execution path.  Native ``b200_*`` methods live in
``native.py`` and are mixed in by ``factory.py``.
"""
from __future__ import annotations

import math
from typing import BinaryIO, List, Tuple

import torch
import torch.nn.functional as F

from .. import model_io as mio
from .base import Layer, Node, pool_out_size


# --------------------------------------------------------------------------- fullc
class FullConnectLayer(Layer):
    """``out = in . W^T + bias`` (reference ``src/layer/fullc_layer-inl.hpp:14-146``)."""
    type_name = "fullc"

    def __init__(self):
        super().__init__()
        self.fullc_gather = 0
        self.fullc_shard = 0      # b200, several ranks: shard the layer by output columns (parallel/tensor_parallel.py)
        self.wmat = self.bias = self.gwmat = self.gbias = None

    def set_param(self, name, val):
        super().set_param(name, val)
        if name == "fullc_gather":
            self.fullc_gather = int(val)
        if name == "fullc_shard":
            self.fullc_shard = int(val)

    def weights(self):
        out = [("wmat", self.wmat, self.gwmat)]
        if self.param.no_bias == 0:
            out.append(("bias", self.bias, self.gbias))
        return out

    def init_connection(self, nodes_in, nodes_out):
        self._check_11(nodes_in, nodes_out, "FullcLayer")
        if not nodes_in[0].is_mat:
            raise ValueError("FullcLayer: input need to be a matrix")
        if self.param.num_hidden <= 0:
            raise ValueError("FullcLayer: must set nhidden correctly")
        nodes_out[0].shape = [nodes_in[0].shape[0], 1, 1, self.param.num_hidden]
        if self.param.num_input_node == 0:
            self.param.num_input_node = nodes_in[0].shape[3]
        elif self.param.num_input_node != nodes_in[0].shape[3]:
            raise ValueError("FullcLayer: input hidden nodes is not consistent")

    def init_model(self):
        p = self.param
        self.wmat = p.rand_init_weight(self.rng, (p.num_hidden, p.num_input_node), p.num_input_node,
                                       p.num_hidden).to(self.device)
        self.bias = torch.full((p.num_hidden,), float(p.init_bias), device=self.device)
        self._alloc_grads()

    def _alloc_grads(self):
        self.gwmat = torch.zeros_like(self.wmat)
        self.gbias = torch.zeros_like(self.bias)

    def save_model(self, fo):
        self.param.save(fo)
        self._save_tensor(fo, self.wmat)
        self._save_tensor(fo, self.bias)

    def load_model(self, fi):
        self.param.load(fi)
        self.wmat = self._load_tensor(fi, 2)
        self.bias = self._load_tensor(fi, 1)
        self._alloc_grads()

    def forward(self, is_train, nodes_in, nodes_out):
        m_in = nodes_in[0].mat()
        out = m_in @ self.wmat.t()
        if self.param.no_bias == 0:
            out = out + self.bias
        nodes_out[0].data = out.view(nodes_out[0].shape)

    def backprop(self, prop_grad, nodes_in, nodes_out):
        m_in = nodes_in[0].mat()
        m_out = nodes_out[0].mat()
        if self.fullc_gather == 0:
            self.gwmat += m_out.t() @ m_in
        if self.param.no_bias == 0:
            self.gbias += m_out.sum(0)
        if prop_grad:
            nodes_in[0].data = (m_out @ self.wmat).view(nodes_in[0].shape)


# --------------------------------------------------------------------------- conv
class ConvolutionLayer(Layer):
    """Grouped cross-correlation (reference ``src/layer/convolution_layer-inl.hpp:12-228``).

    Weight is stored ``(ngroup, nchannel/ngroup, cin/ngroup*kh*kw)`` with the reduction index
    ordered (c, kh, kw); output size ``(H+2p-k)/s+1``.
    """
    type_name = "conv"

    def __init__(self):
        super().__init__()
        self.wmat = self.bias = self.gwmat = self.gbias = None

    def weights(self):
        out = [("wmat", self.wmat, self.gwmat)]
        if self.param.no_bias == 0:
            out.append(("bias", self.bias, self.gbias))
        return out

    def init_connection(self, nodes_in, nodes_out):
        self._check_11(nodes_in, nodes_out, "ConvolutionLayer")
        p = self.param
        b, c, h, w = nodes_in[0].shape
        if c % p.num_group:
            raise ValueError("input channels must divide group size")
        if p.num_channel % p.num_group:
            raise ValueError("output channels must divide group size")
        if p.num_channel <= 0:
            raise ValueError("must set nchannel correctly")
        if p.kernel_height <= 0 or p.kernel_width <= 0:
            raise ValueError("must set kernel_size correctly")
        if p.kernel_width > w or p.kernel_height > h:
            raise ValueError("kernel size exceed input")
        nodes_out[0].shape = [b, p.num_channel, (h + 2 * p.pad_y - p.kernel_height) // p.stride + 1,
                              (w + 2 * p.pad_x - p.kernel_width) // p.stride + 1]
        if p.num_input_channel == 0:
            p.num_input_channel = c
        elif p.num_input_channel != c:
            raise ValueError("ConvolutionLayer: number of input channels is not consistent")

    def init_model(self):
        p = self.param
        shape = (p.num_group, p.num_channel // p.num_group,
                 p.num_input_channel // p.num_group * p.kernel_height * p.kernel_width)
        self.wmat = p.rand_init_weight(self.rng, shape, shape[2], shape[1]).to(self.device)
        self.bias = torch.full((p.num_channel,), float(p.init_bias), device=self.device)
        self.gwmat = torch.zeros_like(self.wmat)
        self.gbias = torch.zeros_like(self.bias)

    def save_model(self, fo):
        self.param.save(fo)
        self._save_tensor(fo, self.wmat)
        self._save_tensor(fo, self.bias)

    def load_model(self, fi):
        self.param.load(fi)
        self.wmat = self._load_tensor(fi, 3)
        self.bias = self._load_tensor(fi, 1)
        self.gwmat = torch.zeros_like(self.wmat)
        self.gbias = torch.zeros_like(self.bias)

    def _w4(self, w=None):
        p = self.param
        w = self.wmat if w is None else w
        return w.reshape(p.num_channel, p.num_input_channel // p.num_group, p.kernel_height, p.kernel_width)

    def forward(self, is_train, nodes_in, nodes_out):
        p = self.param
        x = nodes_in[0].data
        self.state["x"] = x  # input activation is still needed by backprop (node keeps it in the reference)
        y = F.conv2d(x, self._w4(), None if p.no_bias else self.bias, stride=p.stride,
                     padding=(p.pad_y, p.pad_x), groups=p.num_group)
        nodes_out[0].data = y

    def backprop(self, prop_grad, nodes_in, nodes_out):
        p = self.param
        x = nodes_in[0].data
        g = nodes_out[0].data
        if p.no_bias == 0:
            self.gbias += g.sum((0, 2, 3))
        w4 = self._w4()
        gw = torch.nn.grad.conv2d_weight(x, w4.shape, g, stride=p.stride, padding=(p.pad_y, p.pad_x),
                                         groups=p.num_group)
        self.gwmat += gw.reshape(self.gwmat.shape)
        if prop_grad:
            nodes_in[0].data = torch.nn.grad.conv2d_input(x.shape, w4, g, stride=p.stride,
                                                          padding=(p.pad_y, p.pad_x), groups=p.num_group)


# --------------------------------------------------------------------------- pooling
def _pad_for_pool(x, ky, kx, s, oh, ow, value):
    hp, wp = (oh - 1) * s + ky, (ow - 1) * s + kx
    ph, pw = max(hp - x.shape[2], 0), max(wp - x.shape[3], 0)
    if ph or pw:
        x = F.pad(x, (0, pw, 0, ph), value=value)
    return x[:, :, :hp, :wp]


def pool_forward(x, ky, kx, s, mode):
    """mode: 'max' | 'sum' | 'avg' with clipped border windows, avg divides by ky*kx."""
    oh, ow = pool_out_size(x.shape[2], ky, s), pool_out_size(x.shape[3], kx, s)
    if mode == "max":
        xp = _pad_for_pool(x, ky, kx, s, oh, ow, float("-inf"))
        return F.max_pool2d(xp, (ky, kx), s)
    xp = _pad_for_pool(x, ky, kx, s, oh, ow, 0.0)
    y = F.avg_pool2d(xp, (ky, kx), s)
    return y * (ky * kx) if mode == "sum" else y


def pool_backward(x, pooled, gpooled, ky, kx, s, mode):
    """unpool: every input equal to the window max receives the gradient (reference ``unpool``)."""
    b, c, h, w = x.shape
    oh, ow = pooled.shape[2], pooled.shape[3]
    hp, wp = (oh - 1) * s + ky, (ow - 1) * s + kx
    L = oh * ow
    if mode == "max":
        xp = _pad_for_pool(x, ky, kx, s, oh, ow, float("nan"))
        xu = F.unfold(xp, (ky, kx), stride=s).view(b, c, ky * kx, L)
        contrib = (xu == pooled.reshape(b, c, 1, L)).to(x.dtype) * gpooled.reshape(b, c, 1, L)
    else:
        contrib = gpooled.reshape(b, c, 1, L).expand(b, c, ky * kx, L)
        if mode == "avg":
            contrib = contrib * (1.0 / (ky * kx))
    gx = F.fold(contrib.reshape(b, c * ky * kx, L), (hp, wp), (ky, kx), stride=s)
    if hp < h or wp < w:
        gx = F.pad(gx, (0, max(w - wp, 0), 0, max(h - hp, 0)))
    return gx[:, :, :h, :w].contiguous()


class PoolingLayer(Layer):
    """max / sum / avg pooling (reference ``src/layer/pooling_layer-inl.hpp:11-114``)."""
    mode = "max"
    fused_relu = False

    def init_connection(self, nodes_in, nodes_out):
        self._check_11(nodes_in, nodes_out, "PoolingLayer")
        p = self.param
        b, c, h, w = nodes_in[0].shape
        if p.kernel_height <= 0 or p.kernel_width <= 0:
            raise ValueError("must set kernel_size correctly")
        if p.kernel_width > w or p.kernel_height > h:
            raise ValueError("kernel size exceed input")
        nodes_out[0].shape = [b, c, pool_out_size(h, p.kernel_height, p.stride),
                              pool_out_size(w, p.kernel_width, p.stride)]

    def forward(self, is_train, nodes_in, nodes_out):
        p = self.param
        if self.fused_relu:
            nodes_in[0].data = torch.relu(nodes_in[0].data)
        y = pool_forward(nodes_in[0].data, p.kernel_height, p.kernel_width, p.stride, self.mode)
        self.state["pooled"] = y
        nodes_out[0].data = y.clone()

    def backprop(self, prop_grad, nodes_in, nodes_out):
        if not prop_grad:
            return
        p = self.param
        x = nodes_in[0].data
        gx = pool_backward(x, self.state["pooled"], nodes_out[0].data, p.kernel_height, p.kernel_width,
                           p.stride, self.mode)
        if self.fused_relu:
            gx = (x > 0).to(gx.dtype) * gx
        nodes_in[0].data = gx


class MaxPoolingLayer(PoolingLayer):
    type_name = "max_pooling"
    mode = "max"


class SumPoolingLayer(PoolingLayer):
    type_name = "sum_pooling"
    mode = "sum"


class AvgPoolingLayer(PoolingLayer):
    type_name = "avg_pooling"
    mode = "avg"


class ReluMaxPoolingLayer(PoolingLayer):
    type_name = "relu_max_pooling"
    mode = "max"
    fused_relu = True


class InsanityPoolingLayer(PoolingLayer):
    """Max-pool whose source pixels are randomly displaced by one position while training
    (reference ``src/layer/insanity_pooling_layer-inl.hpp:63-94,162-214,253-258``).

    The reference bounds the x displacement in its unpool by the *channel* extent
    (``shape_[srcdim-3]``, ``:166``); the image width is used here.
    """
    type_name = "insanity_max_pooling"
    mode = "max"

    def __init__(self):
        super().__init__()
        self.p_keep = 1.0

    def set_param(self, name, val):
        super().set_param(name, val)
        if name == "keep":
            self.p_keep = float(val)

    def _displaced(self, x, mask):
        b, c, h, w = x.shape
        delta = (1.0 - self.p_keep) / 4.0
        ys = torch.arange(h, device=x.device).view(1, 1, h, 1).expand(b, c, h, w)
        xs = torch.arange(w, device=x.device).view(1, 1, 1, w).expand(b, c, h, w)
        k = self.p_keep
        up = (mask >= k) & (mask < k + delta)
        dn = (mask >= k + delta) & (mask < k + 2 * delta)
        lf = (mask >= k + 2 * delta) & (mask < k + 3 * delta)
        rt = mask >= k + 3 * delta
        ly = torch.where(up, (ys - 1).clamp(min=0), ys)
        ly = torch.where(dn, (ys + 1).clamp(max=h - 1), ly)
        lx = torch.where(lf, (xs - 1).clamp(min=0), xs)
        lx = torch.where(rt, (xs + 1).clamp(max=w - 1), lx)
        return x.reshape(b, c, h * w).gather(2, (ly * w + lx).reshape(b, c, h * w)).view(b, c, h, w)

    def forward(self, is_train, nodes_in, nodes_out):
        p = self.param
        x = nodes_in[0].data
        if is_train:
            mask = torch.rand(x.shape, generator=self.rng).to(x.device)
            self.state["mask"] = mask
            src = self._displaced(x, mask)
        else:
            src = x
        y = pool_forward(src, p.kernel_height, p.kernel_width, p.stride, "max")
        self.state["pooled"] = y
        nodes_out[0].data = y.clone()

    def backprop(self, prop_grad, nodes_in, nodes_out):
        p = self.param
        x = nodes_in[0].data
        src = self._displaced(x, self.state["mask"]) if "mask" in self.state else x
        nodes_in[0].data = pool_backward(src, self.state["pooled"], nodes_out[0].data, p.kernel_height,
                                         p.kernel_width, p.stride, "max")


# --------------------------------------------------------------------------- activations
class ActivationLayer(Layer):
    """In place on the input node, then copied to the output (reference
    ``src/layer/activation_layer-inl.hpp:11-40``, functors ``op.h:27-73``)."""
    op = "relu"

    def init_connection(self, nodes_in, nodes_out):
        self._check_11(nodes_in, nodes_out, "ActivationLayer")
        nodes_out[0].shape = list(nodes_in[0].shape)

    def _fwd(self, x):
        return {"relu": torch.relu, "sigmoid": torch.sigmoid, "tanh": torch.tanh}[self.op](x)

    def _grad(self, a):
        if self.op == "relu":
            return (a > 0).to(a.dtype)
        if self.op == "sigmoid":
            return a * (1.0 - a)
        return 1.0 - a * a

    def forward(self, is_train, nodes_in, nodes_out):
        nodes_in[0].data = self._fwd(nodes_in[0].data)
        nodes_out[0].data = nodes_in[0].data.clone()

    def backprop(self, prop_grad, nodes_in, nodes_out):
        if prop_grad:
            nodes_in[0].data = self._grad(nodes_in[0].data) * nodes_out[0].data


class ReluLayer(ActivationLayer):
    type_name = "relu"
    op = "relu"


class SigmoidLayer(ActivationLayer):
    type_name = "sigmoid"
    op = "sigmoid"


class TanhLayer(ActivationLayer):
    type_name = "tanh"
    op = "tanh"


class XeluLayer(ActivationLayer):
    """leaky unit ``a>0 ? a : a/b`` (reference ``src/layer/xelu_layer-inl.hpp:14-50``)."""
    type_name = "xelu"
    op = "xelu"

    def __init__(self):
        super().__init__()
        self.b = 5.0

    def set_param(self, name, val):
        super().set_param(name, val)
        if name == "b":
            self.b = float(val)

    def _fwd(self, x):
        return torch.where(x > 0, x, x / self.b)

    def _grad(self, a):
        return torch.where(a > 0, torch.ones_like(a), torch.full_like(a, 1.0 / self.b))


class InsanityLayer(Layer):
    """Randomised leaky ReLU (reference ``src/layer/insanity_layer-inl.hpp:13-102``)."""
    type_name = "insanity"

    def __init__(self):
        super().__init__()
        self.lb, self.ub = 5.0, 10.0
        self.step = 0
        self.sat_start = self.sat_end = 0
        self.delta = 0.0
        self.inited = False

    def set_param(self, name, val):
        super().set_param(name, val)
        if name == "lb": self.lb = float(val)
        if name == "ub": self.ub = float(val)
        if name == "calm_start": self.sat_start = int(val)
        if name == "calm_end": self.sat_end = int(val)

    def init_connection(self, nodes_in, nodes_out):
        self._check_11(nodes_in, nodes_out, "InsanityLayer")
        nodes_out[0].shape = list(nodes_in[0].shape)

    def _advance(self):
        """the calm-down schedule of [lb, ub], evaluated once per forward pass (reference :49-60)"""
        if not self.inited:
            self.inited = True
            d = self.ub - (self.ub + self.lb) / 2.0
            span = self.sat_end - self.sat_start
            self.delta = d / span if span != 0 else float("inf")
        if self.sat_start < self.step < self.sat_end:
            self.ub -= self.delta * self.step
            self.lb += self.delta * self.step
            self.step += 1

    def forward(self, is_train, nodes_in, nodes_out):
        self._advance()
        x = nodes_in[0].data
        if is_train:
            mask = torch.rand(x.shape, generator=self.rng).to(x.device) * (self.ub - self.lb) + self.lb
            self.state["mask"] = mask
            y = torch.where(x > 0, x, x / mask)
        else:
            y = torch.where(x > 0, x, x / ((self.lb + self.ub) / 2.0))
        nodes_in[0].data = y
        nodes_out[0].data = y.clone()

    def backprop(self, prop_grad, nodes_in, nodes_out):
        if prop_grad:
            a = nodes_in[0].data
            nodes_in[0].data = torch.where(a > 0, torch.ones_like(a), 1.0 / self.state["mask"]) * nodes_out[0].data


class PReluLayer(Layer):
    """Per-channel learned slope (reference ``src/layer/prelu_layer-inl.hpp:47-173``); the slope is
    visited under tag ``bias`` and checkpointed without a LayerParam header."""
    type_name = "prelu"
    weight_attrs = {"bias": ("slope", "gslope")}

    def __init__(self):
        super().__init__()
        self.init_slope = 0.25
        self.init_random = 0
        self.random = 0.0
        self.channel = 0
        self.slope = self.gslope = None

    def set_param(self, name, val):
        super().set_param(name, val)
        if name == "init_slope": self.init_slope = float(val)
        if name == "random_slope": self.init_random = int(val)
        if name == "random": self.random = float(val)

    def weights(self):
        return [("bias", self.slope, self.gslope)]

    def init_connection(self, nodes_in, nodes_out):
        self._check_11(nodes_in, nodes_out, "PReluLayer")
        nodes_out[0].shape = list(nodes_in[0].shape)
        self.channel = nodes_in[0].shape[3] if nodes_in[0].shape[1] == 1 else nodes_in[0].shape[1]

    def init_model(self):
        if self.init_random == 0:
            self.slope = torch.full((self.channel,), float(self.init_slope), device=self.device)
        else:
            self.slope = (torch.rand((self.channel,), generator=self.rng) * self.init_slope).to(self.device)
        self.gslope = torch.zeros_like(self.slope)

    def save_model(self, fo):
        self._save_tensor(fo, self.slope)

    def load_model(self, fi):
        self.slope = self._load_tensor(fi, 1)
        self.gslope = torch.zeros_like(self.slope)

    def _bshape(self, x):
        return (1, -1, 1, 1) if x.shape[1] != 1 else (1, 1, 1, -1)

    def forward(self, is_train, nodes_in, nodes_out):
        x = nodes_in[0].data
        mask = self.slope.view(self._bshape(x)).expand_as(x)
        if is_train:
            noise = torch.rand(x.shape, generator=self.rng).to(x.device)
            mask = mask * (1 + noise * self.random * 2.0 - self.random)
        mask = mask.clamp(0.0, 1.0)
        self.state["mask"] = mask
        nodes_out[0].data = torch.where(x > 0, x, x * mask)

    def backprop(self, prop_grad, nodes_in, nodes_out):
        x = nodes_in[0].data
        g = nodes_out[0].data
        dims = (0, 2, 3) if x.shape[1] != 1 else (0, 1, 2)
        self.gslope += (torch.where(x > 0, torch.zeros_like(x), x) * g).sum(dims)
        if prop_grad:
            nodes_in[0].data = torch.where(x > 0, torch.ones_like(x), self.state["mask"]) * g


# --------------------------------------------------------------------------- dropout / flatten / bias
class DropoutLayer(Layer):
    """self-loop; ``mask = (U < pkeep)/pkeep`` (reference ``src/layer/dropout_layer-inl.hpp:11-67``)."""
    type_name = "dropout"
    self_loop = True

    def __init__(self):
        super().__init__()
        self.pkeep = 1.0

    def set_param(self, name, val):
        super().set_param(name, val)
        if name == "threshold":
            t = float(val)
            if not (0.0 <= t < 1.0):
                raise ValueError("DropoutLayer: invalid dropout threshold")
            self.pkeep = 1.0 - t

    def init_connection(self, nodes_in, nodes_out):
        self._check_11(nodes_in, nodes_out, "DropoutLayer")
        if nodes_in[0] is not nodes_out[0]:
            raise ValueError("DropoutLayer is an self-loop Layer")

    def forward(self, is_train, nodes_in, nodes_out):
        if is_train:
            x = nodes_out[0].data
            mask = (torch.rand(x.shape, generator=self.rng).to(x.device) < self.pkeep).to(x.dtype) / self.pkeep
            self.state["mask"] = mask
            nodes_out[0].data = x * mask

    def backprop(self, prop_grad, nodes_in, nodes_out):
        if prop_grad:
            nodes_out[0].data = nodes_out[0].data * self.state["mask"]


class FlattenLayer(Layer):
    """reference ``src/layer/flatten_layer-inl.hpp:10-39``."""
    type_name = "flatten"

    def init_connection(self, nodes_in, nodes_out):
        self._check_11(nodes_in, nodes_out, "FlattenLayer")
        b, c, h, w = nodes_in[0].shape
        nodes_out[0].shape = [b, 1, 1, c * h * w]

    def forward(self, is_train, nodes_in, nodes_out):
        nodes_out[0].data = nodes_in[0].data.reshape(nodes_out[0].shape).clone()

    def backprop(self, prop_grad, nodes_in, nodes_out):
        if prop_grad:
            nodes_in[0].data = nodes_out[0].data.reshape(nodes_in[0].shape).clone()


class BiasLayer(Layer):
    """self-loop bias on a matrix node (reference ``src/layer/bias_layer-inl.hpp:13-82``)."""
    type_name = "bias"
    self_loop = True

    def __init__(self):
        super().__init__()
        self.bias = self.gbias = None

    def weights(self):
        return [("bias", self.bias, self.gbias)]

    def init_connection(self, nodes_in, nodes_out):
        self._check_11(nodes_in, nodes_out, "BiasLayer")
        if nodes_in[0] is not nodes_out[0]:
            raise ValueError("BiasLayer is an self-loop Layer")
        if not nodes_in[0].is_mat:
            raise ValueError("BiasLayer: input need to be a matrix")
        if self.param.num_input_node == 0:
            self.param.num_input_node = nodes_in[0].shape[3]

    def init_model(self):
        self.bias = torch.full((self.param.num_input_node,), float(self.param.init_bias), device=self.device)
        self.gbias = torch.zeros_like(self.bias)

    def save_model(self, fo):
        self.param.save(fo)
        self._save_tensor(fo, self.bias)

    def load_model(self, fi):
        self.param.load(fi)
        self.bias = self._load_tensor(fi, 1)
        self.gbias = torch.zeros_like(self.bias)

    def forward(self, is_train, nodes_in, nodes_out):
        nodes_in[0].data = nodes_in[0].data + self.bias.view(1, 1, 1, -1)

    def backprop(self, prop_grad, nodes_in, nodes_out):
        self.gbias += nodes_in[0].mat().sum(0)


# --------------------------------------------------------------------------- lrn / batch norm
def _chpool_sum(t, nsize):
    half = nsize // 2
    c = t.shape[1]
    tp = F.pad(t, (0, 0, 0, 0, half, half))
    out = torch.zeros_like(t)
    for j in range(2 * half + 1):
        out = out + tp[:, j:j + c]
    return out


class LRNLayer(Layer):
    """cross-channel normalisation (reference ``src/layer/lrn_layer-inl.hpp:11-90``)."""
    type_name = "lrn"

    def __init__(self):
        super().__init__()
        self.knorm, self.nsize, self.alpha, self.beta = 1.0, 3, 0.0, 0.0

    def set_param(self, name, val):
        super().set_param(name, val)
        if name == "local_size": self.nsize = int(val)
        if name == "alpha": self.alpha = float(val)
        if name == "beta": self.beta = float(val)
        if name == "knorm": self.knorm = float(val)

    def init_connection(self, nodes_in, nodes_out):
        self._check_11(nodes_in, nodes_out, "LRNLayer")
        nodes_out[0].shape = list(nodes_in[0].shape)

    def forward(self, is_train, nodes_in, nodes_out):
        x = nodes_in[0].data
        salpha = self.alpha / self.nsize
        norm = _chpool_sum(x * x, self.nsize) * salpha + self.knorm
        self.state["norm"] = norm
        nodes_out[0].data = x * norm.pow(-self.beta)

    def backprop(self, prop_grad, nodes_in, nodes_out):
        if not prop_grad:
            return
        x = nodes_in[0].data
        g = nodes_out[0].data
        norm = self.state["norm"]
        salpha = self.alpha / self.nsize
        gx = g * norm.pow(-self.beta)
        gx = gx + (-2.0 * self.beta * salpha) * _chpool_sum(g * x * norm.pow(-self.beta - 1.0), self.nsize) * x
        nodes_in[0].data = gx


class BatchNormLayer(Layer):
    """Batch normalisation that uses mini-batch statistics at test time too
    (reference ``src/layer/batch_norm_layer-inl.hpp``, eps default 1e-10)."""
    type_name = "batch_norm"
    weight_attrs = {"wmat": ("slope", "gslope"), "bias": ("bias", "gbias")}

    def __init__(self):
        super().__init__()
        self.init_slope, self.init_bias, self.eps = 1.0, 0.0, 1e-10
        self.channel = 0
        self.slope = self.gslope = self.bias = self.gbias = None

    def set_param(self, name, val):
        super().set_param(name, val)
        if name == "init_slope": self.init_slope = float(val)
        if name == "init_bias": self.init_bias = float(val)
        if name == "eps": self.eps = float(val)

    def weights(self):
        return [("wmat", self.slope, self.gslope), ("bias", self.bias, self.gbias)]

    def init_connection(self, nodes_in, nodes_out):
        self._check_11(nodes_in, nodes_out, "BNLayer")
        nodes_out[0].shape = list(nodes_in[0].shape)
        self.channel = nodes_in[0].shape[3] if nodes_in[0].shape[1] == 1 else nodes_in[0].shape[1]

    def init_model(self):
        self.slope = torch.full((self.channel,), float(self.init_slope), device=self.device)
        self.bias = torch.full((self.channel,), float(self.init_bias), device=self.device)
        self.gslope = torch.zeros_like(self.slope)
        self.gbias = torch.zeros_like(self.bias)

    def save_model(self, fo):
        self._save_tensor(fo, self.slope)
        self._save_tensor(fo, self.bias)

    def load_model(self, fi):
        self.slope = self._load_tensor(fi, 1)
        self.bias = self._load_tensor(fi, 1)
        self.gslope = torch.zeros_like(self.slope)
        self.gbias = torch.zeros_like(self.bias)

    def _dims(self, x):
        return ((0, 2, 3), (1, -1, 1, 1)) if x.shape[1] != 1 else ((0, 1, 2), (1, 1, 1, -1))

    def forward(self, is_train, nodes_in, nodes_out):
        x = nodes_in[0].data
        dims, bs = self._dims(x)
        scale = 1.0 / x.numel() * self.channel
        mean = scale * x.sum(dims)
        var = scale * ((x - mean.view(bs)) ** 2).sum(dims)
        self.state["mean"], self.state["var"] = mean, var
        if is_train:
            self.state["x"] = x.clone()
            xhat = (x - mean.view(bs)) / torch.sqrt(var + self.eps).view(bs)
            nodes_in[0].data = xhat
            nodes_out[0].data = xhat * self.slope.view(bs) + self.bias.view(bs)
        else:
            istd = 1.0 / torch.sqrt(var + self.eps)
            nodes_out[0].data = (self.slope * istd).view(bs) * x + (self.bias - self.slope * mean * istd).view(bs)

    def backprop(self, prop_grad, nodes_in, nodes_out):
        xhat = nodes_in[0].data
        g = nodes_out[0].data
        x = self.state["x"]
        mean, var = self.state["mean"], self.state["var"]
        dims, bs = self._dims(x)
        scale = 1.0 / x.numel() * self.channel
        xc = x - mean.view(bs)
        gs = g * self.slope.view(bs)
        gvar = (gs * xc * -0.5 * (var + self.eps).pow(-1.5).view(bs)).sum(dims)
        gexp = gs.sum(dims) * (-1.0 / torch.sqrt(var + self.eps))
        gexp = gexp + scale * (-2.0 * xc).sum(dims) * gvar
        self.gslope += (g * xhat).sum(dims)
        self.gbias += g.sum(dims)
        nodes_in[0].data = (gs * (1.0 / torch.sqrt(var + self.eps)).view(bs) + gvar.view(bs) * scale * 2.0 * xc
                            + gexp.view(bs) * scale)


# --------------------------------------------------------------------------- split / concat / fixconn
class SplitLayer(Layer):
    """1 -> n copy, gradients summed (reference ``src/layer/split_layer-inl.hpp:11-45``)."""
    type_name = "split"

    def init_connection(self, nodes_in, nodes_out):
        if len(nodes_in) != 1 or len(nodes_out) < 1:
            raise ValueError("SplitLayer: only support 1-n connection")
        for n in nodes_out:
            n.shape = list(nodes_in[0].shape)

    def forward(self, is_train, nodes_in, nodes_out):
        for n in nodes_out:
            n.data = nodes_in[0].data.clone()

    def backprop(self, prop_grad, nodes_in, nodes_out):
        if prop_grad:
            g = nodes_out[0].data.clone()
            for n in nodes_out[1:]:
                g = g + n.data
            nodes_in[0].data = g


class ConcatLayer(Layer):
    """n -> 1 concatenation along ``dim`` (reference ``src/layer/concat_layer-inl.hpp:11-79``)."""
    type_name = "concat"
    dim = 3

    def init_connection(self, nodes_in, nodes_out):
        if len(nodes_in) <= 1 or len(nodes_out) != 1:
            raise ValueError("Concat layer only support n-1 connection")
        if len(nodes_in) > 4:
            raise ValueError("More than 4 input node is unspported")
        oshape = list(nodes_in[0].shape)
        total = 0
        for n in nodes_in:
            total += n.shape[self.dim]
            for j in range(4):
                if j != self.dim and n.shape[j] != oshape[j]:
                    raise ValueError("Concat shape doesn't match")
        oshape[self.dim] = total
        nodes_out[0].shape = oshape

    def forward(self, is_train, nodes_in, nodes_out):
        nodes_out[0].data = torch.cat([n.data for n in nodes_in], dim=self.dim)

    def backprop(self, prop_grad, nodes_in, nodes_out):
        if prop_grad:
            parts = torch.split(nodes_out[0].data, [n.shape[self.dim] for n in nodes_in], dim=self.dim)
            for n, p in zip(nodes_in, parts):
                n.data = p.contiguous()


class ChConcatLayer(ConcatLayer):
    type_name = "ch_concat"
    dim = 1


class FixConnectLayer(Layer):
    """Fixed sparse matrix loaded from text ``nrow ncol nnz`` + ``x y val`` triples
    (reference ``src/layer/fixconn_layer-inl.hpp:14-94``)."""
    type_name = "fixconn"

    def __init__(self):
        super().__init__()
        self.fname = ""
        self.wmat = None

    def set_param(self, name, val):
        super().set_param(name, val)
        if name == "fixconn_weight":
            self.fname = val

    def init_connection(self, nodes_in, nodes_out):
        self._check_11(nodes_in, nodes_out, "FixConnLayer")
        if not nodes_in[0].is_mat:
            raise ValueError("FixConnLayer: input need to be a matrix")
        if self.param.num_hidden <= 0:
            raise ValueError("FixConnLayer: must set nhidden correctly")
        nodes_out[0].shape = [nodes_in[0].shape[0], 1, 1, self.param.num_hidden]
        if self.param.num_input_node == 0:
            self.param.num_input_node = nodes_in[0].shape[3]
        if not self.fname:
            raise ValueError("FixConnLayer: must specify fixconn_weight")
        with open(self.fname) as f:
            toks = f.read().split()
        nrow, ncol, nnz = int(toks[0]), int(toks[1]), int(toks[2])
        if nrow != self.param.num_hidden or ncol != self.param.num_input_node:
            raise ValueError("FixConnLayer: fixconn_weight shape do not match architecture")
        w = torch.zeros(nrow, ncol)
        for i in range(nnz):
            x, y, v = int(toks[3 + 3 * i]), int(toks[4 + 3 * i]), float(toks[5 + 3 * i])
            w[x, y] = v
        self.wmat = w.to(self.device)

    def to_device(self, device):
        super().to_device(device)
        if self.wmat is not None:
            self.wmat = self.wmat.to(device)

    def forward(self, is_train, nodes_in, nodes_out):
        nodes_out[0].data = (nodes_in[0].mat() @ self.wmat.t()).view(nodes_out[0].shape)

    def backprop(self, prop_grad, nodes_in, nodes_out):
        if prop_grad:
            nodes_in[0].data = (nodes_out[0].mat() @ self.wmat).view(nodes_in[0].shape)


# --------------------------------------------------------------------------- losses
class LossLayerBase(Layer):
    """self-loop loss layer; gradient scaled by ``grad_scale/(batch_size*update_period)`` with the
    *global* batch size (reference ``src/layer/loss/loss_layer_base-inl.hpp:11-133``)."""
    self_loop = True
    allow_sharing = False

    def __init__(self):
        super().__init__()
        self.batch_size = 0
        self.update_period = 1
        self.target = "label"
        self.grad_scale = 1.0
        self.target_index = 0

    def set_param(self, name, val):
        super().set_param(name, val)
        if name == "batch_size": self.batch_size = int(val)
        if name == "update_period": self.update_period = int(val)
        if name == "target": self.target = val
        if name == "grad_scale": self.grad_scale = float(val)

    def init_connection(self, nodes_in, nodes_out):
        self._check_11(nodes_in, nodes_out, "LossLayer")
        if nodes_in[0] is not nodes_out[0]:
            raise ValueError("LossLayer is an self-loop Layer")
        m = self.label_info.name2findex
        if self.target not in m:
            raise ValueError("LossLayer: unknown target=%s" % self.target)
        self.target_index = m[self.target]

    def scale(self) -> float:
        return self.grad_scale / (self.batch_size * self.update_period)

    def _transform(self, x):
        return x

    def _grad(self, p, label):
        raise NotImplementedError

    def forward(self, is_train, nodes_in, nodes_out):
        n = nodes_in[0]
        n.data = self._transform(n.mat()).view(n.shape)

    def backprop(self, prop_grad, nodes_in, nodes_out):
        n = nodes_in[0]
        label = self.label_info.fields[self.target_index].to(n.data.device)
        n.data = (self._grad(n.mat(), label) * self.scale()).view(n.shape)


class SoftmaxLayer(LossLayerBase):
    """reference ``src/layer/loss/softmax_layer-inl.hpp:12-33``."""
    type_name = "softmax"

    def _transform(self, x):
        return torch.softmax(x, dim=1)

    def _grad(self, p, label):
        g = p.clone()
        idx = label[:, 0].long()
        g[torch.arange(p.shape[0], device=p.device), idx] -= 1.0
        return g


class L2LossLayer(LossLayerBase):
    """reference ``src/layer/loss/l2_loss_layer-inl.hpp``."""
    type_name = "l2_loss"

    def _grad(self, p, label):
        if label.shape[1] != p.shape[1]:
            raise ValueError("L2LossLayer: label size mismatch")
        return p - label


class MultiLogisticLayer(LossLayerBase):
    """reference ``src/layer/loss/multi_logistic_layer-inl.hpp``."""
    type_name = "multi_logistic"

    def _transform(self, x):
        return torch.sigmoid(x)

    def _grad(self, p, label):
        if label.shape[1] != p.shape[1]:
            raise ValueError("MultiLogisticLayer: label size mismatch")
        return p - label
