"""Native sm_100a (``b200`` backend) implementations of the hot layers.

Activations are NHWC bf16 (``ops.DevMat``), weights are fp32 masters in the parameter arena
(``parallel/arena.py``) with a bf16 shadow that the tcgen05 GEMM reads through TMA.  Each
class extends the reference-math class from ``zoo.py`` and adds ``b200_*`` methods; the
reference methods stay available as the oracle.  Kernel inventory: SURVEY.md section 2.9.
"""
from __future__ import annotations

import os

import torch

from .. import ops
from ..ops import DevMat
from . import zoo
from .base import Node


def _rows(node: Node) -> int:
    if node.tp_rows:          # stored [global batch][this rank's columns]
        return node.tp_rows
    b, c, h, w = node.shape
    return b if node.is_mat else b * h * w


class _ParamBinding:
    """Arena views handed to a native layer for one tag."""
    __slots__ = ("w", "g", "shadow", "shadow_ld", "rows", "cols", "seg")

    def __init__(self, w, g, shadow, shadow_ld, rows, cols, seg=None):
        self.w, self.g, self.shadow, self.shadow_ld, self.rows, self.cols = w, g, shadow, shadow_ld, rows, cols
        self.seg = seg        # the arena segment (optimiser state, entry of the per-step hyper-parameter table)


# ----------------------------------------------------------------------------- fullc
class FullConnectLayer(zoo.FullConnectLayer):
    """K5/K6: three tcgen05 GEMMs, bias (+ReLU) fused into the forward epilogue."""
    has_b200 = True

    def __init__(self):
        super().__init__()
        self.fuse_relu = False
        self.bind = {}
        self.shard = None     # parallel.exchange.FcShard when trained with fullc_gather on several ranks
        self.tp = None        # parallel.tensor_parallel.TpLayer when sharded by output columns (fullc_shard)

    def b200_shadow(self, tag):
        p = self.param
        if tag == "wmat":
            return (p.num_hidden, p.num_input_node, ops.round_up(p.num_input_node, 8))
        return None

    def b200_forward(self, ex, is_train):
        if self.tp is not None:
            ex.tp.fc_forward(self, is_train)
            return
        p = self.param
        x, y = self.nodes_in[0], self.nodes_out[0]
        b = x.shape[0]
        W = self.bind["wmat"]
        bias = self.bind["bias"].w.data_ptr() if p.no_bias == 0 else None
        flags = ops.GEMM_RELU if self.fuse_relu else 0
        if y.loss_input:
            dptr, ldd = y.f32.data_ptr(), y.f32.shape[1]
        else:
            dptr, ldd = y.act.ptr, y.act.ld
            flags |= ops.GEMM_OUT_BF16
        ops.gemm(x.act.ptr, x.act.ld, 0, 0, W.shadow.data_ptr(), W.shadow_ld, 0, 0, dptr, ldd, 0, bias, 0,
                 b, p.num_hidden, p.num_input_node, flags=flags)

    def b200_backprop(self, ex, prop_grad):
        if self.tp is not None:
            ex.tp.fc_backward(self, prop_grad)
            return
        p = self.param
        x, y = self.nodes_in[0], self.nodes_out[0]
        b = x.shape[0]
        W = self.bind["wmat"]
        gy = y.grad
        if self.shard is not None and ex.exchange is not None:
            # fullc_gather: backward-data locally, weight gradient / update row-sharded over the ranks
            if prop_grad:
                ops.gemm(gy.ptr, gy.ld, 0, 0, W.shadow.data_ptr(), W.shadow_ld, 0, 1, x.grad.ptr, x.grad.ld, 0, None, 0,
                         b, p.num_input_node, p.num_hidden, flags=ops.GEMM_OUT_BF16)
            ex.exchange.fc_backward(self, ex.need_update)
            return
        # the updater fused into the weight-gradient epilogue (this step updates, nothing accumulates, one GPU):
        # the gradient matrix never reaches memory; backward-data goes first because it reads the old weights
        fuse = ex.fuse_now and W.seg is not None and W.seg.fused
        if not fuse:
            # gW (nout x nin) += gy^T . x   : A = gy stored [K=b][M=nout], B = x stored [K=b][N=nin]
            ops.gemm(gy.ptr, gy.ld, 0, 1, x.act.ptr, x.act.ld, 0, 1, W.g.data_ptr(), p.num_input_node, 0, None, 0,
                     p.num_hidden, p.num_input_node, b, flags=ops.GEMM_BETA1 if ex.accumulate_grads else 0)
        if p.no_bias == 0:
            ops.check(ops.lib().cxn_colsum_bf16(gy.ptr, b, p.num_hidden, gy.ld, self.bind["bias"].g.data_ptr(),
                                                 1 if ex.accumulate_grads else 0, ops.stream()), "colsum")
        if prop_grad:
            # gx (b x nin) = gy (b x nout) . W (nout x nin) : B = W stored [K=nout][N=nin]
            ops.gemm(gy.ptr, gy.ld, 0, 0, W.shadow.data_ptr(), W.shadow_ld, 0, 1, x.grad.ptr, x.grad.ld, 0, None, 0,
                     b, p.num_input_node, p.num_hidden, flags=ops.GEMM_OUT_BF16)
        if fuse:
            a, sg = ex.arena, W.seg
            ops.check(ops.lib().cxn_gemm_wgrad_update(
                gy.ptr, gy.ld, x.act.ptr, x.act.ld, p.num_hidden, p.num_input_node, b,
                W.w.data_ptr(), a.m.data_ptr() + 4 * sg.start, (a.m2.data_ptr() + 4 * sg.start) if a.m2 is not None else None,
                W.shadow.data_ptr(), p.num_input_node, W.shadow_ld, a.hyper_ptr(sg), a.kind, ops.stream()),
                "cxn_gemm_wgrad_update")


# ----------------------------------------------------------------------------- conv
class ConvolutionLayer(zoo.ConvolutionLayer):
    """K1-K4 on tensor cores.

    Default path: implicit GEMM (``csrc/kernels/conv_igemm_sm100.cu``) - no im2col buffer; the
    patch operand is gathered by cp.async producer warps, the filter / dy operand arrives by
    TMA, groups are one batched launch.  Forward fuses bias (+ReLU); backward-data of stride-1
    convolutions is the same kernel run on dy with flipped filters; backward-weight is the
    split-K variant with fp32 atomics.  Shapes the implicit kernel does not cover (channel
    counts that are not a multiple of 8, strided backward-data) use an explicit NHWC patch
    gather + the plain tcgen05 GEMM.

    Internal weight layout is ``(g, oc/g, kh, kw, ic/g)`` (channels innermost, matching the
    NHWC gather; a 3-channel input is padded to 4); ``weight_to_internal`` /
    ``weight_from_internal`` convert to and from the checkpoint layout ``(g, oc/g, ic/g*kh*kw)``.
    """
    has_b200 = True

    def __init__(self):
        super().__init__()
        self.fuse_relu = False
        self.bind = {}
        self.col = None
        self.wflip = None
        self.cin_pad = 0        # 4 when a <=4-channel input is stored with a channel pitch of 4
        self.pair = False       # cin_pad input read as [H][Wp/2] super-pixels of 8 channels (two pixels each)
        self.mask_relu_in = False   # backward-data also applies the mask of the ReLU that produced the input node
        self.span = False       # cin_pad input read through im2col TMA as overlapping 16-pixel (64-channel) positions
        self.igemm_fwd = False
        self.igemm_dgrad = False
        self.bias_grad_fused = False   # the pooling layer behind this one accumulates the bias gradient (_NativePool.bias_of)
        self.dgrad_phases = False      # strided backward-data as stride*stride implicit-GEMM problems (cxn_conv_dgrad_strided)
        self.phase_scratch = None
        self.cpad = 0           # internal channels per group, padded up to a multiple of 64 (see b200_plan)
        self.s2d = False        # stride-4 first layer over a space-to-depth input node (see b200_plan)
        self.s2d_mask = None

    def b200_supported(self):
        p = self.param
        return (p.num_channel // p.num_group) % 8 == 0

    def b200_plan(self, ex, input_is_node0: bool, sole_consumer: bool):
        """decide kernel paths (called before buffers and the parameter arena are laid out)."""
        p = self.param
        g = p.num_group
        icg = p.num_input_channel // g
        small = p.kernel_height <= 16 and p.kernel_width <= 16
        if icg % 8 == 0 and small and (p.kernel_height * p.kernel_width * icg) // 8 <= 1024:
            self.igemm_fwd = True
        elif (g == 1 and p.num_input_channel <= 4 and input_is_node0 and sole_consumer and small):
            self.igemm_fwd = True
            self.cin_pad = 4
            # Pixel pairs: with an even stride and no horizontal padding every filter row starts on an even
            # pixel, so the image (stored with an even row pitch) reads as super-pixels of two pixels = 8
            # channels = one 16-byte granule, filter width ceil(kw/2), stride_x / 2.  Half the gather
            # instructions of the 8-byte (one pixel) granule form; the odd filter column, if any, is a
            # structural zero of the internal weight layout.
            self.pair = p.stride % 2 == 0 and p.pad_x == 0 and os.environ.get("CXN_CONV_PAIR", "1") != "0"
            # Spans: a filter row of the pitch-4 image is one contiguous run, so "16 pixels x 4 channels" is a
            # 64-channel position that starts every `stride` pixels (positions overlap in memory - a tensor map
            # does not mind).  That makes the layer a kh x 1 convolution over 64 channels which im2col-mode TMA
            # fetches without a single gather instruction; filter columns kw..15 are structural zeros.
            self.span = (self.pair and p.kernel_width <= 16 and p.kernel_height <= 16
                         and os.environ.get("CXN_CONV_SPAN", "1") != "0")
            # Space to depth: with stride 4 and no padding the layer is a STRIDE-1 convolution with a
            # ceil(kh/4) x ceil(kw/4) filter over an image of 4 x 4 pixel blocks = 64 channels (4 per pixel).  A patch
            # row is then one ALIGNED 128-byte line (a span starts every 32 bytes and straddles two lines, which is
            # what bounds the span route: half the L2 -> SM rate of the aligned layers), the reduction shrinks from
            # kh * 64 to ceil(kh/4) * ceil(kw/4) * 64, and the layer takes the ordinary im2col-TMA kernels.  Filter
            # taps past the kh x kw window are structural zeros (masked out of the weight gradient).
            b_, c_, h_, w_ = self.nodes_in[0].shape
            k3h, k3w = (p.kernel_height + 3) // 4, (p.kernel_width + 3) // 4
            self.s2d = (p.stride == 4 and p.pad_x == 0 and p.pad_y == 0 and k3h * k3w * 8 <= 1024
                        and (h_ + 3) // 4 - k3h + 1 == self.nodes_out[0].shape[2]
                        and (w_ + 3) // 4 - k3w + 1 == self.nodes_out[0].shape[3]
                        and p.num_channel % 8 == 0 and os.environ.get("CXN_CONV_S2D", "1") != "0")
            if self.s2d:
                self.pair = self.span = False
        ocg = p.num_channel // g
        pad_ok = p.kernel_height - 1 - p.pad_y >= 0 and p.kernel_width - 1 - p.pad_x >= 0
        self.igemm_dgrad = (self.igemm_fwd and p.stride == 1 and ocg % 8 == 0 and pad_ok
                            and (p.kernel_height * p.kernel_width * ocg) // 8 <= 1024)
        # Channel-padded groups: the im2col-mode TMA route fetches 64-channel chunks, and a layer with e.g. 48
        # channels per group (AlexNet conv2) would otherwise build its patch operand with gather warps, which bound
        # its forward and weight-gradient kernels.  Instead the filter's internal layout pads every tap to 64 channels
        # with structural zeros and each group runs as a one-group convolution over a channel window that starts at
        # the group's first channel: what the window reads beyond the group (the next group's channels, or zeros past
        # the tensor) meets zero weights.  A third more MMA work, no gather instructions.  Stride-1 layers whose
        # backward-data is the implicit kernel (it works on the unpadded filter) only.
        b, c, h, w = self.nodes_in[0].shape
        oh, ow = self.nodes_out[0].shape[2], self.nodes_out[0].shape[3]
        cp = ops.round_up(icg, 64)
        if (self.igemm_fwd and self.igemm_dgrad and not self.cin_pad and icg % 64 != 0 and icg % 8 == 0 and 2 * cp <= 3 * icg
                and (p.kernel_height * p.kernel_width * cp) // 8 <= 1024 and c % 8 == 0
                and max(p.pad_x, p.pad_y, p.kernel_width - 1 - p.pad_x, p.kernel_height - 1 - p.pad_y) <= 127
                and (w + 2 * p.pad_x - p.kernel_width) // p.stride + 1 == ow
                and (h + 2 * p.pad_y - p.kernel_height) // p.stride + 1 == oh
                and os.environ.get("CXN_CONV_CPAD", "1") != "0"):
            self.cpad = cp
        if self.s2d:
            return ("s2d", 4)
        if self.pair:
            return (self.cin_pad, ops.round_up(self.nodes_in[0].shape[3], 2))
        return self.cin_pad

    def _icg(self):
        """channels per group of the internal filter layout"""
        p = self.param
        if self.s2d:
            return 64
        return self.cin_pad or self.cpad or p.num_input_channel // p.num_group

    def _k3(self):
        """filter extent in 4 x 4 pixel blocks (space-to-depth layers)"""
        return (self.param.kernel_height + 3) // 4, (self.param.kernel_width + 3) // 4

    def _group_window(self, x, gi):
        """(pointer, visible channels, row stride, image stride) of group gi's channel window (cpad layers)"""
        p = self.param
        b, c, h, w = x.shape
        cin = p.num_input_channel // p.num_group
        ld = x.act.ld
        return x.act.ptr + 2 * gi * cin, p.num_input_channel - gi * cin, w * ld, h * w * ld

    def _kw2(self):
        return (self.param.kernel_width + 1) // 2

    def _geom(self):
        p = self.param
        g = p.num_group
        icg = self._icg()
        kg = icg * p.kernel_height * p.kernel_width
        if self.pair:
            kg = 8 * p.kernel_height * self._kw2()
        if self.span:
            kg = 64 * p.kernel_height
        if self.s2d:
            k3h, k3w = self._k3()
            kg = 64 * k3h * k3w
        return g, icg, p.num_channel // g, kg, ops.round_up(kg, 8)

    def _gather_args(self, x):
        """(H, W, pixel pitch, channels per group, granule, KW, stride, pad_x) as the implicit-GEMM kernel sees the input"""
        p = self.param
        b, c, h, w = x.shape
        if self.pair:
            return (h, x.wpitch // 2, 8, 8, 8, self._kw2(), p.stride | ((p.stride // 2) << 16), 0)
        return (h, w, x.act.ld, self._icg(), 4 if self.cin_pad else 8, p.kernel_width, p.stride, p.pad_x)

    def weight_to_internal(self, tag, w):
        if tag != "wmat":
            return w
        p = self.param
        g, icg, ocg, kg, _ = self._geom()
        cin = p.num_input_channel // g
        if self.s2d:
            # (oc, c, kh, kw) -> (oc, ky, kx, dy, dx, c4) with kh = 4 ky + dy, kw = 4 kx + dx; zeros elsewhere
            k3h, k3w = self._k3()
            w4 = w.reshape(ocg, cin, p.kernel_height, p.kernel_width)
            w4 = torch.nn.functional.pad(w4, (0, 4 * k3w - p.kernel_width, 0, 4 * k3h - p.kernel_height, 0, 4 - cin))
            return w4.reshape(ocg, 4, k3h, 4, k3w, 4).permute(0, 2, 4, 3, 5, 1).reshape(1, ocg, kg)
        w5 = w.reshape(g, ocg, cin, p.kernel_height, p.kernel_width).permute(0, 1, 3, 4, 2)
        if icg != cin:
            w5 = torch.nn.functional.pad(w5, (0, icg - cin))
        if self.span:
            w5 = torch.nn.functional.pad(w5, (0, 0, 0, 16 - p.kernel_width))                # zero filter columns kw..15
        elif self.pair:
            w5 = torch.nn.functional.pad(w5, (0, 0, 0, 2 * self._kw2() - p.kernel_width))   # zero odd filter column
        return w5.reshape(g, ocg, kg)

    def weight_from_internal(self, tag, w):
        if tag != "wmat":
            return w
        p = self.param
        g, icg, ocg, kg, _ = self._geom()
        cin = p.num_input_channel // g
        if self.s2d:
            k3h, k3w = self._k3()
            w4 = w.reshape(ocg, k3h, k3w, 4, 4, 4).permute(0, 5, 1, 3, 2, 4).reshape(ocg, 4, 4 * k3h, 4 * k3w)
            return w4[:, :cin, :p.kernel_height, :p.kernel_width].reshape(1, ocg, cin * p.kernel_height * p.kernel_width)
        kw = 16 if self.span else (2 * self._kw2() if self.pair else p.kernel_width)
        w5 = w.reshape(g, ocg, p.kernel_height, kw, icg)[:, :, :, :p.kernel_width, :cin]
        return w5.permute(0, 1, 4, 2, 3).reshape(g, ocg, cin * p.kernel_height * p.kernel_width)

    def b200_shadow(self, tag):
        if tag == "wmat":
            g, icg, ocg, kg, kpad = self._geom()
            return (g * ocg, kg, kpad)
        return None

    def b200_setup(self, ex):
        p = self.param
        g, icg, ocg, kg, kpad = self._geom()
        y = self.nodes_out[0]
        L = ops.lib()
        if (self.igemm_fwd and not self.igemm_dgrad and not self.cin_pad and p.stride > 1 and ocg % 8 == 0 and icg % 8 == 0
                and y.grad.ld % 8 == 0 and os.environ.get("CXN_CONV_DGRAD_PHASES", "1") != "0"
                and L.cxn_conv_dgrad_strided_ok(p.kernel_height, p.kernel_width, p.stride, p.pad_y, p.pad_x) == 1):
            # strided layer: backward-data by stride phases, no patch buffer
            self.dgrad_phases = True
            n = int(L.cxn_conv_dgrad_strided_scratch(g, ocg, icg, p.kernel_height, p.kernel_width, p.stride))
            self.phase_scratch = torch.zeros((max(n, 8),), dtype=torch.bfloat16, device=ex.device)
        if self.s2d:
            # structural zeros of the internal layout, as a 0 / 1 row: applied to the weight gradient
            ones = torch.ones((1, ocg, (p.num_input_channel // g) * p.kernel_height * p.kernel_width))
            self.s2d_mask = self.weight_to_internal("wmat", ones)[0, 0].contiguous().to(ex.device, torch.float32)
        elif (not self.igemm_fwd or not self.igemm_dgrad) and not (self.dgrad_phases and self.igemm_fwd):
            max_rows = ex.max_batch * y.shape[2] * y.shape[3]
            self.col = torch.empty((g * max_rows * kpad,), dtype=torch.bfloat16, device=ex.device)
        self.ev_dy, self.ev_wg = torch.cuda.Event(), torch.cuda.Event()
        if self.igemm_dgrad:
            self.wflip_ld = ops.round_up(p.kernel_height * p.kernel_width * ocg, 8)
            self.wflip = torch.zeros((g * (p.num_input_channel // g if self.cpad else icg) * self.wflip_ld,),
                                     dtype=torch.bfloat16, device=ex.device)

    def _im2col_args(self, x, y, rows, kpad):
        p = self.param
        b, c, h, w = x.shape
        return (b, h, w, c, p.num_group, p.kernel_height, p.kernel_width, p.stride, p.pad_y, p.pad_x,
                y.shape[2], y.shape[3], kpad, rows * kpad)

    def b200_forward(self, ex, is_train):
        p = self.param
        x, y = self.nodes_in[0], self.nodes_out[0]
        g, icg, ocg, kg, kpad = self._geom()
        rows = _rows(y)
        L = ops.lib()
        W = self.bind["wmat"]
        bias = self.bind["bias"].w.data_ptr() if p.no_bias == 0 else None
        b, c, h, w = x.shape
        if self.s2d:
            k3h, k3w = self._k3()
            ops.check(L.cxn_conv_fprop(x.act.ptr, b, (h + 3) // 4, (w + 3) // 4, 64, 64, 1, 8, W.shadow.data_ptr(), kpad, ocg,
                                       k3h, k3w, 1, 0, 0, y.shape[2], y.shape[3], y.act.ptr, y.act.ld, bias,
                                       1 if self.fuse_relu else 0, ops.stream()), "conv_fprop")
            return
        if self.span:
            wp = x.wpitch
            ops.check(L.cxn_conv_fprop_view(x.act.ptr, b, h, y.shape[3], p.stride * 4, 64, 1, W.shadow.data_ptr(), kpad, ocg,
                                            p.kernel_height, 1, p.stride | (1 << 16), p.pad_y, 0, y.shape[2], y.shape[3],
                                            y.act.ptr, y.act.ld, bias, 1 if self.fuse_relu else 0, 64, wp * 4, h * wp * 4,
                                            ops.stream()), "conv_fprop_view")
            return
        if self.cpad:
            # all groups in one launch: group g reads the 64-channel chunks of the window that starts at channel g * cin
            ptr, cdim, rstride, istride = self._group_window(x, 0)
            ops.check(L.cxn_conv_view_group_stride(p.num_input_channel // g), "group_stride")
            ops.check(L.cxn_conv_fprop_view(ptr, b, h, w, x.act.ld, self.cpad, g, W.shadow.data_ptr(), kpad, ocg,
                                            p.kernel_height, p.kernel_width, p.stride, p.pad_y, p.pad_x, y.shape[2], y.shape[3],
                                            y.act.ptr, y.act.ld, bias, 1 if self.fuse_relu else 0, cdim, rstride, istride,
                                            ops.stream()), "conv_fprop_view")
            return
        if self.igemm_fwd:
            gh, gw, gcs, gcg, gran, gkw, gstride, gpad_x = self._gather_args(x)
            ops.check(L.cxn_conv_fprop(x.act.ptr, b, gh, gw, gcs, gcg, g, gran,
                                       W.shadow.data_ptr(), kpad, ocg, p.kernel_height, gkw, gstride,
                                       p.pad_y, gpad_x, y.shape[2], y.shape[3], y.act.ptr, y.act.ld, bias,
                                       1 if self.fuse_relu else 0, ops.stream()), "conv_fprop")
            return
        ops.check(L.cxn_im2col_nhwc(x.act.ptr, self.col.data_ptr(), *self._im2col_args(x, y, rows, kpad),
                                    ops.stream()), "im2col")
        flags = ops.GEMM_OUT_BF16 | (ops.GEMM_RELU if self.fuse_relu else 0)
        ops.gemm(self.col.data_ptr(), kpad, rows * kpad, 0, W.shadow.data_ptr(), kpad, ocg * kpad, 0,
                 y.act.ptr, y.act.ld, ocg, bias, ocg, rows, ocg, kg, batch=g, flags=flags)

    def _weight_gradient(self, ex):
        """accumulates into the arena gradient, which the updater leaves zeroed"""
        p = self.param
        x, y = self.nodes_in[0], self.nodes_out[0]
        g, icg, ocg, kg, kpad = self._geom()
        rows = _rows(y)
        gy = y.grad
        L = ops.lib()
        W = self.bind["wmat"]
        b, c, h, w = x.shape
        if self.s2d:
            k3h, k3w = self._k3()
            ops.check(L.cxn_conv_wgrad(x.act.ptr, b, (h + 3) // 4, (w + 3) // 4, 64, 64, 1, 8, gy.ptr, gy.ld, ocg, k3h, k3w, 1,
                                       0, 0, y.shape[2], y.shape[3], W.g.data_ptr(), kg, 0, ops.stream()), "conv_wgrad")
            # filter taps outside the kh x kw window met real pixels: they are structural zeros and stay zero
            ops.check(L.cxn_mask_cols_f32(W.g.data_ptr(), ocg, kg, kg, self.s2d_mask.data_ptr(), ops.stream()), "mask_cols")
        elif self.span:
            wp = x.wpitch
            ops.check(L.cxn_conv_wgrad_view(x.act.ptr, b, h, y.shape[3], p.stride * 4, 64, 1, gy.ptr, gy.ld, ocg,
                                            p.kernel_height, 1, p.stride | (1 << 16), p.pad_y, 0, y.shape[2], y.shape[3],
                                            W.g.data_ptr(), kg, 0, 64, 4 * p.kernel_width, 64, wp * 4, h * wp * 4,
                                            ops.stream()), "conv_wgrad_view")
        elif self.cpad:
            # the padding channels of every tap are structural zeros: their gradient (real values of the next group)
            # must not accumulate (mask period / valid)
            cin = p.num_input_channel // g
            ptr, cdim, rstride, istride = self._group_window(x, 0)
            ops.check(L.cxn_conv_view_group_stride(cin), "group_stride")
            ops.check(L.cxn_conv_wgrad_view(ptr, b, h, w, x.act.ld, self.cpad, g, gy.ptr, gy.ld, ocg, p.kernel_height,
                                            p.kernel_width, p.stride, p.pad_y, p.pad_x, y.shape[2], y.shape[3],
                                            W.g.data_ptr(), kg, 0, self.cpad, cin, cdim, rstride, istride, ops.stream()),
                      "conv_wgrad_view")
        elif self.igemm_fwd:
            gh, gw, gcs, gcg, gran, gkw, gstride, gpad_x = self._gather_args(x)
            # pair mode: the odd filter column is a structural zero - its gradient (real pixels!) must not accumulate
            mask = (8 * gkw, 4 * p.kernel_width) if self.pair and p.kernel_width % 2 else (0, 0)
            ops.check(L.cxn_conv_wgrad_masked(x.act.ptr, b, gh, gw, gcs, gcg, g, gran, gy.ptr, gy.ld,
                                              ocg, p.kernel_height, gkw, gstride, p.pad_y, gpad_x, y.shape[2],
                                              y.shape[3], W.g.data_ptr(), kg, 0, mask[0], mask[1], ops.stream()),
                      "conv_wgrad")
        else:
            kb = (rows + 63) // 64
            tiles = ((ocg + 127) // 128) * ((kg + 127) // 128) * g
            splits = max(1, min(kb // 4 if kb >= 8 else 1, (2 * ex.num_sms + tiles - 1) // tiles))
            # (col still holds the forward patches)
            ops.gemm(gy.ptr, gy.ld, ocg, 1, self.col.data_ptr(), kpad, rows * kpad, 1, W.g.data_ptr(), kg, ocg * kg,
                     None, 0, ocg, kg, rows, batch=g, splits=splits, flags=ops.GEMM_ATOMIC)

    def _dgrad_tiles(self, x) -> int:
        """CTAs the stride-1 backward-data launch occupies (one 128-row tile each; the kernels are persistent, one CTA per SM)"""
        p = self.param
        b, c, h, w = x.shape
        icg = p.num_input_channel // p.num_group
        bnmax = 256 if icg > 128 else 128
        return ((b * h * w + 127) // 128) * ((icg + bnmax - 1) // bnmax) * p.num_group

    def b200_backprop(self, ex, prop_grad):
        p = self.param
        x, y = self.nodes_in[0], self.nodes_out[0]
        g, icg, ocg, kg, kpad = self._geom()
        rows = _rows(y)
        gy = y.grad
        L = ops.lib()
        W = self.bind["wmat"]
        b, c, h, w = x.shape
        if p.no_bias == 0 and not self.bias_grad_fused:
            ops.check(L.cxn_colsum_bf16(gy.ptr, rows, p.num_channel, gy.ld, self.bind["bias"].g.data_ptr(),
                                        1 if ex.accumulate_grads else 0, ops.stream()), "colsum")
        # Weight gradient.  When the backward-data problem of this layer fills only part of the GPU (small per-GPU
        # batches: conv3 of AlexNet at 32 images is 43 tiles on 148 SMs) the two independent kernels run SIDE BY SIDE:
        # the weight gradient goes to a side stream with its grid capped to the SMs backward-data leaves free.
        side = None
        if (prop_grad and self.igemm_dgrad and self.igemm_fwd and not self.span and ex.wgrad_stream is not None):
            tiles = self._dgrad_tiles(x)
            if tiles * 4 <= ex.num_sms * 3:
                side = ex.num_sms - tiles
        if side is not None:
            main = torch.cuda.current_stream(ex.device)
            self.ev_dy.record(main)
            ex.wgrad_stream.wait_event(self.ev_dy)
            with torch.cuda.stream(ex.wgrad_stream):
                prev = L.cxn_conv_wgrad_grid_cap(side)
                try:
                    self._weight_gradient(ex)
                finally:
                    L.cxn_conv_wgrad_grid_cap(prev)
                self.ev_wg.record(ex.wgrad_stream)
        else:
            self._weight_gradient(ex)
        if not prop_grad:
            return
        if self.igemm_dgrad:
            kh, kw = p.kernel_height, p.kernel_width
            if self.cpad:
                icg = p.num_input_channel // g      # backward-data works on the real channels of the group
                ops.check(L.cxn_conv_flip_weights_pitch(W.shadow.data_ptr(), self.wflip.data_ptr(), g, ocg, icg, kh, kw, kpad,
                                                        self.wflip_ld, self.cpad, ops.stream()), "flip_weights")
            else:
                ops.check(L.cxn_conv_flip_weights(W.shadow.data_ptr(), self.wflip.data_ptr(), g, ocg, icg, kh, kw, kpad,
                                                  self.wflip_ld, ops.stream()), "flip_weights")
            ops.check(L.cxn_conv_fprop_masked(gy.ptr, b, y.shape[2], y.shape[3], gy.ld, ocg, g, 8, self.wflip.data_ptr(),
                                              self.wflip_ld, icg, kh, kw, 1, kh - 1 - p.pad_y, kw - 1 - p.pad_x, h, w,
                                              x.grad.ptr, x.grad.ld, x.act.ptr if self.mask_relu_in else None,
                                              ops.stream()), "conv_dgrad")
            if side is not None:
                torch.cuda.current_stream(ex.device).wait_event(self.ev_wg)      # join: the layer is done when both are
            return
        if self.dgrad_phases:
            ops.check(L.cxn_conv_dgrad_strided(gy.ptr, b, y.shape[2], y.shape[3], gy.ld, ocg, g, W.shadow.data_ptr(), kpad, icg,
                                               p.kernel_height, p.kernel_width, p.stride, p.pad_y, p.pad_x, h, w, x.grad.ptr,
                                               x.grad.ld, None, self.phase_scratch.data_ptr(), ops.stream()), "conv_dgrad_strided")
            return
        # strided backward-data the explicit way: dcol = gy . W, then scatter
        if self.col is None:      # (space-to-depth first layer: only an input-gradient request ever gets here)
            self.col = torch.empty((g * ex.max_batch * y.shape[2] * y.shape[3] * kpad,), dtype=torch.bfloat16, device=ex.device)
        ops.gemm(gy.ptr, gy.ld, ocg, 0, W.shadow.data_ptr(), kpad, ocg * kpad, 1, self.col.data_ptr(), kpad,
                 rows * kpad, None, 0, rows, kg, ocg, batch=g, flags=ops.GEMM_OUT_BF16)
        if self.s2d:
            k3h, k3w = self._k3()
            ops.check(L.cxn_col2im_nhwc(self.col.data_ptr(), x.grad.ptr, b, (h + 3) // 4, (w + 3) // 4, 64, 1, k3h, k3w, 1, 0, 0,
                                        y.shape[2], y.shape[3], kpad, rows * kpad, ops.stream()), "col2im")
            return
        # pair mode: (kh, kw2, 8) columns are (kh, 2*kw2, 4) in the pixel domain of the pitched image
        cw, ckw = (x.wpitch, 16 if self.span else 2 * self._kw2()) if self.pair else (w, p.kernel_width)
        ops.check(L.cxn_col2im_nhwc(self.col.data_ptr(), x.grad.ptr, b, h, cw, x.act.ld, p.num_group, p.kernel_height,
                                    ckw, p.stride, p.pad_y, p.pad_x, y.shape[2], y.shape[3], kpad,
                                    rows * kpad, ops.stream()), "col2im")


# ----------------------------------------------------------------------------- activations
class _NativeAct:
    has_b200 = True
    act_op = ops.ACT_RELU
    alias_ok = False    # out node may share the in node's buffers (sign-preserving activations)
    fused = False       # forward folded into the producing GEMM epilogue
    skip_bwd = False    # backward mask folded into the consuming pooling layer

    def _b(self):
        return getattr(self, "b", 1.0)

    def b200_forward(self, ex, is_train):
        x, y = self.nodes_in[0], self.nodes_out[0]
        if self.fused:
            return
        rows = _rows(x)
        L = ops.lib()
        ops.check(L.cxn_act_fwd(x.act.ptr, x.act.ptr, rows, x.act.cols, x.act.ld, self.act_op, self._b(),
                                ops.stream()), "act_fwd")
        if y.act.ptr != x.act.ptr:
            ops.check(L.cxn_copy2d(x.act.ptr, x.act.ld, y.act.ptr, y.act.ld, rows, x.act.cols, 0, 0, ops.stream()),
                      "copy2d")

    def b200_backprop(self, ex, prop_grad):
        if not prop_grad or self.skip_bwd:
            return
        x, y = self.nodes_in[0], self.nodes_out[0]
        rows = _rows(x)
        ops.check(ops.lib().cxn_act_bwd(x.act.ptr, y.grad.ptr, x.grad.ptr, rows, x.act.cols, x.act.ld, self.act_op,
                                        self._b(), ops.stream()), "act_bwd")


class ReluLayer(_NativeAct, zoo.ReluLayer):
    act_op = ops.ACT_RELU
    alias_ok = True


class SigmoidLayer(_NativeAct, zoo.SigmoidLayer):
    act_op = ops.ACT_SIGMOID


class TanhLayer(_NativeAct, zoo.TanhLayer):
    act_op = ops.ACT_TANH


class XeluLayer(_NativeAct, zoo.XeluLayer):
    act_op = ops.ACT_XELU
    alias_ok = True


# ----------------------------------------------------------------------------- pooling / lrn
class _NativePool:
    has_b200 = True
    pool_mode = ops.POOL_MAX
    mask_relu = False   # backward also applies the mask of the ReLU that produced the input node
    bias_of = None      # convolution whose (ReLU-masked) output this layer pools: its bias gradient = column sums of the
                        # input gradient computed here, accumulated by the same kernel (set by the executor)

    def _args(self, bwd=False):
        p = self.param
        x, y = self.nodes_in[0], self.nodes_out[0]
        b, c, h, w = x.shape
        relu = self.fused_relu or (bwd and self.mask_relu)
        return (b, h, w, c, p.kernel_height, p.kernel_width, p.stride, y.shape[2], y.shape[3], self.pool_mode,
                1 if relu else 0)

    def b200_forward(self, ex, is_train):
        x, y = self.nodes_in[0], self.nodes_out[0]
        ops.check(ops.lib().cxn_pool_fwd(x.act.ptr, y.act.ptr, *self._args(), ops.stream()), "pool_fwd")
        if self.fused_relu:  # reference applies relu in place on the input node
            ops.check(ops.lib().cxn_act_fwd(x.act.ptr, x.act.ptr, _rows(x), x.act.cols, x.act.ld, ops.ACT_RELU, 1.0,
                                            ops.stream()), "act_fwd")

    def b200_backprop(self, ex, prop_grad):
        if not prop_grad:
            return
        x, y = self.nodes_in[0], self.nodes_out[0]
        L = ops.lib()
        conv = self.bias_of
        if conv is not None:
            gb = conv.bind["bias"].g
            if not ex.accumulate_grads:
                ops.check(L.cxn_fill_f32(gb.data_ptr(), gb.numel(), 0.0, ops.stream()), "fill")
            rc = L.cxn_pool_bwd_bias(x.act.ptr, y.act.ptr, y.grad.ptr, x.grad.ptr, gb.data_ptr(), *self._args(True), ops.stream())
            if rc == 0:
                return
            if rc != -8:
                ops.check(rc, "pool_bwd_bias")
        ops.check(L.cxn_pool_bwd(x.act.ptr, y.act.ptr, y.grad.ptr, x.grad.ptr, *self._args(True), ops.stream()),
                  "pool_bwd")
        if conv is not None:      # shape outside the fused kernel: the separate column sum after all
            ops.check(L.cxn_colsum_bf16(x.grad.ptr, _rows(x), conv.param.num_channel, x.grad.ld, conv.bind["bias"].g.data_ptr(),
                                        1, ops.stream()), "colsum")


class MaxPoolingLayer(_NativePool, zoo.MaxPoolingLayer):
    pool_mode = ops.POOL_MAX


class SumPoolingLayer(_NativePool, zoo.SumPoolingLayer):
    pool_mode = ops.POOL_SUM


class AvgPoolingLayer(_NativePool, zoo.AvgPoolingLayer):
    pool_mode = ops.POOL_AVG


class ReluMaxPoolingLayer(_NativePool, zoo.ReluMaxPoolingLayer):
    pool_mode = ops.POOL_MAX


class LRNLayer(zoo.LRNLayer):
    """K10/K11: one warp per pixel, channel vector in shared memory; norm is recomputed in
    backward instead of being stored."""
    has_b200 = True

    def b200_supported(self):
        return not self.nodes_in[0].is_mat and self.nodes_in[0].shape[1] * 4 * 3 * 4 <= 48 * 1024

    def b200_forward(self, ex, is_train):
        x, y = self.nodes_in[0], self.nodes_out[0]
        ops.check(ops.lib().cxn_lrn_fwd(x.act.ptr, y.act.ptr, _rows(x), x.shape[1], self.nsize, self.alpha,
                                        self.beta, self.knorm, ops.stream()), "lrn_fwd")

    def b200_backprop(self, ex, prop_grad):
        if not prop_grad:
            return
        x, y = self.nodes_in[0], self.nodes_out[0]
        ops.check(ops.lib().cxn_lrn_bwd(x.act.ptr, y.grad.ptr, x.grad.ptr, _rows(x), x.shape[1], self.nsize,
                                        self.alpha, self.beta, self.knorm, ops.stream()), "lrn_bwd")


# ----------------------------------------------------------------------------- flatten / dropout
class FlattenLayer(zoo.FlattenLayer):
    """NHWC -> (c,h,w)-ordered rows, so fullc weights keep the checkpoint's column order."""
    has_b200 = True

    def _run(self, src, dst, to_chw):
        x, y = self.nodes_in[0], self.nodes_out[0]
        b, c, h, w = x.shape
        if x.is_mat:
            ops.check(ops.lib().cxn_copy2d(src.ptr, src.ld, dst.ptr, dst.ld, b, w, 0, 0, ops.stream()), "copy2d")
            return
        hw = h * w
        if to_chw:
            ops.check(ops.lib().cxn_bf16_hwc_chw(src.ptr, dst.ptr, b, c, hw, hw * c, dst.ld, c, hw, 1, ops.stream()),
                      "flatten")
        else:
            ops.check(ops.lib().cxn_bf16_hwc_chw(src.ptr, dst.ptr, b, c, hw, src.ld, hw * c, c, hw, 0, ops.stream()),
                      "unflatten")

    def b200_forward(self, ex, is_train):
        self._run(self.nodes_in[0].act, self.nodes_out[0].act, True)

    def b200_backprop(self, ex, prop_grad):
        if prop_grad:
            self._run(self.nodes_out[0].grad, self.nodes_in[0].grad, False)


class DropoutLayer(zoo.DropoutLayer):
    """K12: Philox mask regenerated from (seed, device step counter); nothing is stored."""
    has_b200 = True

    def _apply(self, ex, m):
        n = self.nodes_out[0]
        seed = (ex.seed * 1000003 + self.layer_index * 7919 + 12345) & 0xFFFFFFFFFFFFFFFF
        ops.check(ops.lib().cxn_dropout(m.ptr, m.ptr, _rows(n), m.cols, m.ld, self.pkeep, seed,
                                        ex.step_counter.data_ptr(), ops.stream()), "dropout")

    def b200_forward(self, ex, is_train):
        if is_train:
            self._apply(ex, self.nodes_out[0].act)

    def b200_backprop(self, ex, prop_grad):
        if prop_grad:
            self._apply(ex, self.nodes_out[0].grad)


# ----------------------------------------------------------------------------- losses
class _NativeLoss:
    """K13: softmax / sigmoid in place on the fp32 loss node, gradient on device."""
    has_b200 = True
    loss_kind = 0       # 0 softmax (class-index label), 1 dense label
    transform = "none"

    def b200_forward(self, ex, is_train):
        n = self.nodes_in[0]
        b, cols = n.shape[0], n.shape[3]
        L = ops.lib()
        if self.transform == "softmax":
            ops.check(L.cxn_softmax_rows(n.f32.data_ptr(), b, cols, n.f32.shape[1], ops.stream()), "softmax")
        elif self.transform == "sigmoid":
            ops.check(L.cxn_sigmoid_rows(n.f32.data_ptr(), b, cols, n.f32.shape[1], ops.stream()), "sigmoid")

    def b200_backprop(self, ex, prop_grad):
        n = self.nodes_in[0]
        b, cols = n.shape[0], n.shape[3]
        label = self.label_info.fields[self.target_index]
        if self.loss_kind == 1 and label.shape[1] != cols:
            raise ValueError("%s: label size mismatch" % self.type_name)
        ops.check(ops.lib().cxn_loss_grad(n.f32.data_ptr(), n.f32.shape[1], label.data_ptr(), label.stride(0),
                                          n.grad.ptr, n.grad.ld, b, cols, self.scale(), self.loss_kind, ops.stream()),
                  "loss_grad")


class SoftmaxLayer(_NativeLoss, zoo.SoftmaxLayer):
    loss_kind = 0
    transform = "softmax"


class L2LossLayer(_NativeLoss, zoo.L2LossLayer):
    loss_kind = 1


class MultiLogisticLayer(_NativeLoss, zoo.MultiLogisticLayer):
    loss_kind = 1
    transform = "sigmoid"


# ----------------------------------------------------------------------------- bias / prelu / batch norm
def _mat_view(node, m):
    """(rows, cols, ld) of a node's activation or gradient matrix: channels are the columns"""
    return _rows(node), m.cols, m.ld


class BiasLayer(zoo.BiasLayer):
    """self-loop bias on a matrix node; gradient = column sums (csrc/kernels/misc_layers.cu)."""
    has_b200 = True

    def __init__(self):
        super().__init__()
        self.bind = {}

    def b200_forward(self, ex, is_train):
        x = self.nodes_in[0]
        rows, cols, ld = _mat_view(x, x.act)
        ops.check(ops.lib().cxn_bias_add(x.act.ptr, x.act.ptr, rows, cols, ld, self.bind["bias"].w.data_ptr(), ops.stream()),
                  "bias_add")

    def b200_backprop(self, ex, prop_grad):
        x = self.nodes_in[0]
        rows, cols, ld = _mat_view(x, x.grad)
        ops.check(ops.lib().cxn_colsum_bf16(x.grad.ptr, rows, cols, ld, self.bind["bias"].g.data_ptr(),
                                             1 if ex.accumulate_grads else 0, ops.stream()), "colsum")


class PReluLayer(zoo.PReluLayer):
    """per-channel learned slope; the training-time slope noise is a Philox stream regenerated in backward"""
    has_b200 = True

    def __init__(self):
        super().__init__()
        self.bind = {}

    def _seed(self, ex):
        return (ex.seed * 1000003 + self.layer_index * 7919 + 4242) & 0xFFFFFFFFFFFFFFFF

    def b200_forward(self, ex, is_train):
        x, y = self.nodes_in[0], self.nodes_out[0]
        rows, cols, ld = _mat_view(x, x.act)
        self._noise = float(self.random) if is_train else 0.0
        ops.check(ops.lib().cxn_prelu_fwd(x.act.ptr, y.act.ptr, rows, cols, ld, y.act.ld, self.bind["bias"].w.data_ptr(),
                                          self._noise, self._seed(ex), ex.step_counter.data_ptr(), ops.stream()), "prelu_fwd")

    def b200_backprop(self, ex, prop_grad):
        x, y = self.nodes_in[0], self.nodes_out[0]
        rows, cols, ld = _mat_view(x, x.act)
        b = self.bind["bias"]
        if not ex.accumulate_grads:
            ops.check(ops.lib().cxn_fill_f32(b.g.data_ptr(), b.g.numel(), 0.0, ops.stream()), "fill")
        ops.check(ops.lib().cxn_prelu_bwd(x.act.ptr, y.grad.ptr, x.grad.ptr if prop_grad else None, rows, cols, ld, y.grad.ld,
                                          b.w.data_ptr(), b.g.data_ptr(), getattr(self, "_noise", 0.0), self._seed(ex),
                                          ex.step_counter.data_ptr(), ops.stream()), "prelu_bwd")


class BatchNormLayer(zoo.BatchNormLayer):
    """mini-batch statistics in training and at test time (as the reference); one column reduction per
    pass plus element-wise kernels with per-channel coefficients.  Unlike the reference the input node
    keeps its value (the reference overwrites it with the normalised activations); xhat lives in a
    private buffer."""
    has_b200 = True

    def __init__(self):
        super().__init__()
        self.bind = {}

    def b200_setup(self, ex):
        x = self.nodes_in[0]
        C = x.act.cols
        self.xhat = DevMat(x.act.t.shape[0], C, x.act.ld, device=ex.device)
        self.scratch = torch.zeros(5 * C, dtype=torch.float32, device=ex.device)
        self.mean = torch.zeros(C, dtype=torch.float32, device=ex.device)
        self.istd = torch.zeros(C, dtype=torch.float32, device=ex.device)

    def b200_forward(self, ex, is_train):
        x, y = self.nodes_in[0], self.nodes_out[0]
        rows, C, ld = _mat_view(x, x.act)
        ops.check(ops.lib().cxn_batchnorm_fwd(x.act.ptr, y.act.ptr, self.xhat.ptr if is_train else None, rows, C, ld, y.act.ld,
                                              self.xhat.ld, self.bind["wmat"].w.data_ptr(), self.bind["bias"].w.data_ptr(),
                                              float(self.eps), self.scratch.data_ptr(), self.mean.data_ptr(),
                                              self.istd.data_ptr(), ops.stream()), "batchnorm_fwd")

    def b200_backprop(self, ex, prop_grad):
        x, y = self.nodes_in[0], self.nodes_out[0]
        rows, C, ld = _mat_view(x, x.act)
        gs, gb = self.bind["wmat"].g, self.bind["bias"].g
        if not ex.accumulate_grads:
            ops.check(ops.lib().cxn_fill_f32(gs.data_ptr(), C, 0.0, ops.stream()), "fill")
            ops.check(ops.lib().cxn_fill_f32(gb.data_ptr(), C, 0.0, ops.stream()), "fill")
        ops.check(ops.lib().cxn_batchnorm_bwd(y.grad.ptr, self.xhat.ptr, x.grad.ptr, rows, C, y.grad.ld, self.xhat.ld, x.grad.ld,
                                              self.bind["wmat"].w.data_ptr(), self.istd.data_ptr(), gs.data_ptr(),
                                              gb.data_ptr(), self.scratch.data_ptr(), 1 if prop_grad else 0, ops.stream()),
                  "batchnorm_bwd")


# ----------------------------------------------------------------------------- split / concat
class SplitLayer(zoo.SplitLayer):
    """1 -> n copies forward, sum of the n gradients backward (strided bf16 copies)"""
    has_b200 = True

    def b200_forward(self, ex, is_train):
        x = self.nodes_in[0]
        rows, cols, ld = _mat_view(x, x.act)
        for n in self.nodes_out:
            ops.check(ops.lib().cxn_copy2d(x.act.ptr, ld, n.act.ptr, n.act.ld, rows, cols, 0, 0, ops.stream()), "copy2d")

    def b200_backprop(self, ex, prop_grad):
        if not prop_grad:
            return
        x = self.nodes_in[0]
        rows, cols, ld = _mat_view(x, x.grad)
        for k, n in enumerate(self.nodes_out):
            ops.check(ops.lib().cxn_copy2d(n.grad.ptr, n.grad.ld, x.grad.ptr, ld, rows, cols, 0, 1 if k else 0, ops.stream()),
                      "copy2d")


class _NativeConcat:
    """concatenation along the innermost stored dimension: the columns of matrix nodes (`concat`) or
    the channels of NHWC nodes (`ch_concat`) - in both cases column ranges of the [rows][C] matrices"""
    has_b200 = True

    def b200_supported(self):
        mats = all(n.is_mat for n in self.nodes_in)
        return mats if self.dim == 3 else not any(n.is_mat for n in self.nodes_in)

    def _copy(self, grad: bool):
        y = self.nodes_out[0]
        ym = y.grad if grad else y.act
        rows = _rows(y)
        off = 0
        for n in self.nodes_in:
            m = n.grad if grad else n.act
            yptr = ym.ptr + off * 2
            if grad:
                ops.check(ops.lib().cxn_copy2d(yptr, ym.ld, m.ptr, m.ld, rows, m.cols, 0, 0, ops.stream()), "copy2d")
            else:
                ops.check(ops.lib().cxn_copy2d(m.ptr, m.ld, yptr, ym.ld, rows, m.cols, 0, 0, ops.stream()), "copy2d")
            off += m.cols

    def b200_forward(self, ex, is_train):
        self._copy(False)

    def b200_backprop(self, ex, prop_grad):
        if prop_grad:
            self._copy(True)


class ConcatLayer(_NativeConcat, zoo.ConcatLayer):
    pass


class ChConcatLayer(_NativeConcat, zoo.ChConcatLayer):
    pass


class InsanityPoolingLayer(_NativePool, zoo.InsanityPoolingLayer):
    """max pooling over a randomly displaced copy of the input (training only): one gather kernel writes the
    displaced copy, the ordinary pooling kernels run on it forward and backward.  As in the reference the
    backward pass hands the gradient to the displaced positions' slots, not back to their sources."""
    pool_mode = ops.POOL_MAX

    def b200_setup(self, ex):
        x = self.nodes_in[0]
        self.src = DevMat(x.act.t.shape[0], x.act.cols, x.act.ld, device=ex.device)
        self._displaced = False

    def _seed(self, ex):
        return (ex.seed * 1000003 + self.layer_index * 7919 + 31337) & 0xFFFFFFFFFFFFFFFF

    def b200_forward(self, ex, is_train):
        x, y = self.nodes_in[0], self.nodes_out[0]
        b, c, h, w = x.shape
        self._displaced = bool(is_train)
        src = x.act
        if is_train:
            ops.check(ops.lib().cxn_insanity_displace(x.act.ptr, self.src.ptr, b, h, w, c, x.act.ld, float(self.p_keep),
                                                      self._seed(ex), ex.step_counter.data_ptr(), ops.stream()),
                      "insanity_displace")
            src = self.src
        ops.check(ops.lib().cxn_pool_fwd(src.ptr, y.act.ptr, *self._args(), ops.stream()), "pool_fwd")

    def b200_backprop(self, ex, prop_grad):
        if not prop_grad:
            return
        x, y = self.nodes_in[0], self.nodes_out[0]
        src = self.src if self._displaced else x.act
        ops.check(ops.lib().cxn_pool_bwd(src.ptr, y.act.ptr, y.grad.ptr, x.grad.ptr, *self._args(True), ops.stream()),
                  "pool_bwd")


# ----------------------------------------------------------------------------- insanity / fixconn
class InsanityLayer(zoo.InsanityLayer):
    """randomised leaky ReLU: the per-element divisor is a Philox stream regenerated in backward.  Unlike
    the reference the input node keeps its value (the reference overwrites it with the output)."""
    has_b200 = True

    @property
    def graph_unsafe(self):
        # [lb, ub] move on the host while the calm-down schedule is active: a captured step would freeze them
        return self.sat_end > self.sat_start

    def _seed(self, ex):
        return (ex.seed * 1000003 + self.layer_index * 7919 + 977) & 0xFFFFFFFFFFFFFFFF

    def b200_forward(self, ex, is_train):
        self._advance()
        x, y = self.nodes_in[0], self.nodes_out[0]
        rows, cols, ld = _mat_view(x, x.act)
        self._random = 1 if is_train else 0
        ops.check(ops.lib().cxn_insanity_fwd(x.act.ptr, y.act.ptr, rows, cols, ld, y.act.ld, float(self.lb), float(self.ub),
                                             self._random, self._seed(ex), ex.step_counter.data_ptr(), ops.stream()),
                  "insanity_fwd")

    def b200_backprop(self, ex, prop_grad):
        if not prop_grad:
            return
        x, y = self.nodes_in[0], self.nodes_out[0]
        rows, cols, ld = _mat_view(x, x.act)
        ops.check(ops.lib().cxn_insanity_bwd(x.act.ptr, y.grad.ptr, x.grad.ptr, rows, cols, ld, y.grad.ld, float(self.lb),
                                             float(self.ub), getattr(self, "_random", 0), self._seed(ex),
                                             ex.step_counter.data_ptr(), ops.stream()), "insanity_bwd")


class FixConnectLayer(zoo.FixConnectLayer):
    """fixed (file-given, untrained) connection matrix: two tcgen05 GEMMs on a bf16 copy of the weights"""
    has_b200 = True

    def b200_setup(self, ex):
        nout, nin = self.wmat.shape
        self.w_ld = ops.round_up(nin, 8)
        self.w_bf16 = torch.zeros((nout, self.w_ld), dtype=torch.bfloat16, device=ex.device)
        self.w_bf16[:, :nin] = self.wmat.to(ex.device)

    def b200_forward(self, ex, is_train):
        x, y = self.nodes_in[0], self.nodes_out[0]
        b = x.shape[0]
        nout, nin = self.wmat.shape
        if y.loss_input:
            dptr, ldd, flags = y.f32.data_ptr(), y.f32.shape[1], 0
        else:
            dptr, ldd, flags = y.act.ptr, y.act.ld, ops.GEMM_OUT_BF16
        ops.gemm(x.act.ptr, x.act.ld, 0, 0, self.w_bf16.data_ptr(), self.w_ld, 0, 0, dptr, ldd, 0, None, 0, b, nout, nin,
                 flags=flags)

    def b200_backprop(self, ex, prop_grad):
        if not prop_grad:
            return
        x, y = self.nodes_in[0], self.nodes_out[0]
        b = x.shape[0]
        nout, nin = self.wmat.shape
        ops.gemm(y.grad.ptr, y.grad.ld, 0, 0, self.w_bf16.data_ptr(), self.w_ld, 0, 1, x.grad.ptr, x.grad.ld, 0, None, 0,
                 b, nin, nout, flags=ops.GEMM_OUT_BF16)

