"""Implicit-GEMM convolution kernels (csrc/kernels/conv_igemm_sm100.cu) against torch conv2d in
fp64 on the same bf16-rounded operands (the fp32 path of the oracle library has TF32 switched off in
tests/conftest.py as well).

Tolerances: products of bf16 operands are exact in fp32 and both sides accumulate in fp32, so what is left is the
rounding of the bf16 OUTPUT (half an ulp = 2^-9 of the value; bounded here by 2^-7 of the largest output) and, for
the fp32 weight gradient, the order of ~10^4..10^5 fp32 additions (bounded by 2e-3 of the largest entry)."""
TOL_BF16_OUT = 2.0 ** -7
TOL_F32_SUM = 2e-3
import os

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _bf(t):
    return t.to(torch.bfloat16)


def _case(N, C, H, W, OC, G, KH, KW, stride, pad, relu=True, c_pitch=None):
    from cxxnet_b200 import ops
    L = ops.lib()
    torch.manual_seed(N + C + H + OC)
    Cs = c_pitch or C
    icg_int = (Cs if c_pitch else C) // G          # channels per group as laid out in K
    ocg = OC // G
    x = _bf(torch.randn(N, H, W, C, device="cuda"))
    xs = torch.zeros(N, H, W, Cs, dtype=torch.bfloat16, device="cuda")
    xs[..., :C] = x
    w = _bf(torch.randn(OC, C // G, KH, KW, device="cuda") * (1.0 / (C // G * KH * KW) ** 0.5))
    bias = torch.randn(OC, device="cuda")
    OH, OW = (H + 2 * pad - KH) // stride + 1, (W + 2 * pad - KW) // stride + 1
    # internal weight layout [G][ocg][kh][kw][icg_int], row pitch padded to 8
    kg = KH * KW * icg_int
    kpad = ops.round_up(kg, 8)
    wi = torch.zeros(G, ocg, KH, KW, icg_int, dtype=torch.bfloat16, device="cuda")
    wi[..., :C // G] = w.view(G, ocg, C // G, KH, KW).permute(0, 1, 3, 4, 2)
    wsh = torch.zeros(G * ocg, kpad, dtype=torch.bfloat16, device="cuda")
    wsh[:, :kg] = wi.reshape(G * ocg, kg)
    gran = 4 if c_pitch == 4 else 8
    # ---- forward
    y = torch.zeros(N * OH * OW, OC, dtype=torch.bfloat16, device="cuda")
    ops.check(L.cxn_conv_fprop(xs.data_ptr(), N, H, W, Cs, icg_int, G, gran, wsh.data_ptr(), kpad, ocg, KH, KW, stride,
                               pad, pad, OH, OW, y.data_ptr(), OC, bias.data_ptr(), 1 if relu else 0, ops.stream()),
              "fprop")
    # oracle in fp64 (a plain direct convolution; cuDNN's fp32 algorithms - FFT / Winograd tiles - carry errors of
    # their own around 1e-3 that would otherwise be charged to the kernel under test)
    xr = x.double().permute(0, 3, 1, 2)
    ref = F.conv2d(xr, w.double(), bias.double(), stride=stride, padding=pad, groups=G).float()
    if relu:
        ref = torch.relu(ref)
    got = y.float().view(N, OH, OW, OC).permute(0, 3, 1, 2)
    err = float((got - ref).abs().max())
    assert err <= TOL_BF16_OUT * max(1.0, float(ref.abs().max())), "fprop err %g" % err
    # ---- weight gradient
    dy = _bf(torch.randn(N, OH, OW, OC, device="cuda") * 0.1)
    dw = torch.zeros(G * ocg, kg, dtype=torch.float32, device="cuda")
    ops.check(L.cxn_conv_wgrad(xs.data_ptr(), N, H, W, Cs, icg_int, G, gran, dy.data_ptr(), OC, ocg, KH, KW, stride,
                               pad, pad, OH, OW, dw.data_ptr(), kg, 0, ops.stream()), "wgrad")
    dwr = torch.nn.grad.conv2d_weight(xr, w.shape, dy.double().permute(0, 3, 1, 2), stride=stride, padding=pad,
                                      groups=G).float()
    gotw = dw.view(G, ocg, KH, KW, icg_int)[..., :C // G].permute(0, 1, 4, 2, 3).reshape(OC, C // G, KH, KW)
    errw = float((gotw - dwr).abs().max())
    assert errw <= TOL_F32_SUM * max(1.0, float(dwr.abs().max())), "wgrad err %g (max %g)" % (errw, float(dwr.abs().max()))
    if c_pitch:
        assert float(dw.view(G, ocg, KH, KW, icg_int)[..., C // G:].abs().max()) == 0.0
    # ---- backward data (stride 1 only: forward kernel on dy with flipped filters)
    if stride == 1 and not c_pitch:
        icg = C // G
        kf = KH * KW * ocg
        kfpad = ops.round_up(kf, 8)
        wt = torch.zeros(G * icg, kfpad, dtype=torch.bfloat16, device="cuda")
        ops.check(L.cxn_conv_flip_weights(wsh.data_ptr(), wt.data_ptr(), G, ocg, icg, KH, KW, kpad, kfpad, ops.stream()),
                  "flip")
        dx = torch.zeros(N * H * W, C, dtype=torch.bfloat16, device="cuda")
        ops.check(L.cxn_conv_fprop(dy.data_ptr(), N, OH, OW, OC, ocg, G, 8, wt.data_ptr(), kfpad, icg, KH, KW, 1,
                                   KH - 1 - pad, KW - 1 - pad, H, W, dx.data_ptr(), C, None, 0, ops.stream()), "dgrad")
        dxr = torch.nn.grad.conv2d_input(xr.shape, w.double(), dy.double().permute(0, 3, 1, 2), stride=1, padding=pad,
                                         groups=G).float()
        gotx = dx.float().view(N, H, W, C).permute(0, 3, 1, 2)
        errx = float((gotx - dxr).abs().max())
        assert errx <= TOL_BF16_OUT * max(1.0, float(dxr.abs().max())), "dgrad err %g" % errx
    torch.cuda.synchronize()


def test_conv_rgb_stride4_pitch4():
    _case(4, 3, 67, 67, 96, 1, 11, 11, 4, 0, c_pitch=4)          # AlexNet conv1 shape family


def test_conv_grouped_5x5():
    _case(3, 96, 27, 27, 256, 2, 5, 5, 1, 2)                    # AlexNet conv2


def test_conv_3x3_wide():
    _case(5, 256, 13, 13, 384, 1, 3, 3, 1, 1)                   # conv3
    _case(5, 384, 13, 13, 384, 2, 3, 3, 1, 1)                   # conv4
    _case(5, 384, 13, 13, 256, 2, 3, 3, 1, 1, relu=False)       # conv5


def test_conv_odd_shapes():
    _case(2, 8, 9, 11, 24, 1, 3, 2, 2, 1)
    _case(7, 16, 14, 14, 48, 2, 1, 1, 1, 0)
    _case(1, 48, 21, 21, 96, 1, 3, 3, 1, 1)                      # kaggle_bowl-like


def test_conv_many_tiles_persistent():
    _case(64, 32, 32, 32, 64, 1, 3, 3, 1, 1)                    # 65536 rows: 512 m-tiles over 148 SMs


def test_conv_flat_route_over_the_input_grid():
    """cxn_conv_tune(12, 1): stride-1 forward problems without padding and with few output columns run over the
    input's pixel grid - a tap is a constant row offset, a tile's patch rows one tiled-mode TMA box - and the rows
    outside the output are discarded: same results as the im2col route, on shapes with row tails, groups, two
    64-channel chunks per tap and 256-row tiles"""
    from cxxnet_b200 import ops
    L = ops.lib()
    L.cxn_conv_tune(12, 1)
    try:
        n0 = L.cxn_conv_flat_launches()
        _case(3, 64, 57, 57, 96, 1, 3, 3, 1, 0)                    # space-to-depth conv1's shape class
        assert L.cxn_conv_flat_launches() == n0 + 1
        _case(2, 128, 30, 29, 48, 2, 2, 2, 1, 0, relu=False)       # groups, 2 x 2 filter
        _case(40, 128, 40, 40, 64, 1, 3, 3, 1, 0)                  # two chunks per tap; enough rows for 256-row tiles
        _case(1, 64, 31, 33, 24, 1, 1, 1, 1, 0)                    # 1 x 1: the grid IS the output
        assert L.cxn_conv_flat_launches() == n0 + 4
        _case(2, 64, 12, 12, 32, 1, 3, 3, 1, 0)                    # 44 % of the rows would be discarded: im2col route
        _case(2, 64, 20, 20, 32, 1, 3, 3, 1, 1)                    # padding: im2col route
        assert L.cxn_conv_flat_launches() == n0 + 4
    finally:
        L.cxn_conv_tune(12, -1)


@pytest.mark.parametrize("atma", [1, 0])
def test_conv_im2col_tma_and_gather_routes_agree_with_reference(atma):
    """channels-per-group % 64 == 0 takes the im2col-mode TMA route by default; the same problems through the
    gather warps (cxn_conv_tune key 6) must give the same answers.  Strides, asymmetric kernels, padding
    larger than one, groups, row counts that are not a multiple of the tile."""
    from cxxnet_b200 import ops
    L = ops.lib()
    L.cxn_conv_tune(6, atma)
    try:
        _case(3, 64, 17, 19, 48, 1, 3, 3, 2, 1)                  # stride 2, odd extents, M tail
        _case(2, 128, 12, 12, 64, 2, 5, 3, 1, 2)                 # groups of 64 channels, 5x3 filter, pad 2
        _case(1, 192, 7, 7, 32, 1, 1, 1, 1, 0)                   # 1x1, 49 rows only
        _case(9, 64, 20, 20, 128, 1, 4, 4, 4, 0, relu=False)     # stride = kernel (no overlap)
        _case(40, 64, 13, 13, 96, 1, 3, 3, 1, 1)                 # enough rows for the row-paired (256-row) tiles
    finally:
        L.cxn_conv_tune(6, -1)


def test_conv_cluster_pairs_share_the_filter_tile():
    """cxn_conv_tune(7, 2): forward layers on the im2col-TMA route run as clusters of two CTAs that fetch half of
    the filter tile each and multicast it to both.  Odd and even tile counts, groups, two tile widths, the
    row-paired tiles, and the backward-data use of the same kernel."""
    from cxxnet_b200 import ops
    L = ops.lib()
    L.cxn_conv_tune(7, 2)
    try:
        _case(5, 256, 13, 13, 384, 1, 3, 3, 1, 1)                # conv3: 7 m-tiles (odd) x 2 n-tiles
        _case(5, 384, 13, 13, 384, 2, 3, 3, 1, 1)                # conv4: groups
        _case(6, 64, 16, 16, 128, 1, 3, 3, 1, 1)                 # 12 m-tiles (even), one n-tile
        _case(40, 64, 13, 13, 96, 1, 3, 3, 1, 1)                 # row-paired (256-row) tiles
        _case(64, 64, 32, 32, 64, 1, 3, 3, 1, 1)                 # many pairs per cluster (persistent loop)
    finally:
        L.cxn_conv_tune(7, 0)


@pytest.mark.parametrize("mode", [1, 2])
def test_conv_halo_tiles_shift_the_taps_through_one_patch(mode):
    """cxn_conv_tune(9, mode): stride-1 problems with <= 64 output columns per group load one padded-width patch per
    tile and 64-channel chunk and read every filter tap as a row-shifted view of it (mode 1: two 128-row halves per
    tile, 2: one).  Forward and backward-data uses, groups, image heights that do not divide by the tile."""
    from cxxnet_b200 import ops
    L = ops.lib()
    L.cxn_conv_tune(9, mode)
    try:
        _case(3, 96, 27, 27, 256, 2, 5, 5, 1, 2)                 # AlexNet conv2: its backward-data is the target (128 -> 48)
        _case(2, 128, 27, 27, 96, 2, 5, 5, 1, 2)                 # forward with 64-channel groups and 48 outputs
        _case(3, 64, 13, 13, 48, 1, 3, 3, 1, 1)                  # 3x3, Wp = 15: eight output rows per half
        _case(2, 64, 9, 11, 32, 1, 3, 2, 1, 1, relu=False)       # asymmetric filter, odd extents
        _case(5, 128, 20, 20, 64, 1, 3, 3, 1, 1)                 # 64 output columns, two channel chunks
    finally:
        L.cxn_conv_tune(9, 0)


def test_conv_cta_pairs_split_the_filter_tile():
    """cxn_conv_tune(10, 2): forward problems on the im2col-TMA route run as CTA pairs (tcgen05 cta_group::2): one
    256-row tile per pair, every CTA stages its own 128 output positions and half of the filter tile.  Odd tile counts
    (the second CTA of the last pair is all padding), groups, both accumulator widths, tile widths that are not a
    multiple of 32, many tiles per pair (persistent loop), and the backward-data use of the same kernel."""
    from cxxnet_b200 import ops
    L = ops.lib()
    L.cxn_conv_tune(10, 2)
    try:
        _case(5, 256, 13, 13, 384, 1, 3, 3, 1, 1)                # conv3: 845 rows = 3.3 pair tiles, 2 n-tiles of 192
        _case(5, 384, 13, 13, 384, 2, 3, 3, 1, 1)                # conv4: groups, bn = 192
        _case(5, 384, 13, 13, 256, 2, 3, 3, 1, 1, relu=False)    # conv5: bn = 128
        _case(6, 64, 16, 16, 128, 1, 3, 3, 1, 1)                 # exactly 6 pair tiles
        _case(3, 64, 17, 19, 48, 1, 3, 3, 2, 1)                  # stride 2, bn = 48
        _case(64, 64, 32, 32, 64, 1, 3, 3, 1, 1)                 # 256 pair tiles on 74 pairs
    finally:
        L.cxn_conv_tune(10, 0)


@pytest.mark.parametrize("C,OC,G,H,W,k,s,pad", [(16, 32, 1, 13, 13, 3, 2, 1), (32, 64, 2, 16, 14, 5, 2, 2), (8, 16, 1, 9, 9, 2, 2, 0),
                                                (64, 64, 1, 12, 12, 4, 2, 1), (16, 16, 1, 11, 10, 2, 3, 0),
                                                (24, 48, 1, 23, 23, 7, 4, 3), (64, 128, 1, 28, 28, 3, 2, 0)])
def test_conv_strided_backward_data_by_stride_phases(C, OC, G, H, W, k, s, pad):
    """cxn_conv_dgrad_strided: backward-data of a strided convolution as stride x stride stride-1 implicit-GEMM problems
    written through a strided view of dx (no patch buffer, no inserted zeros), against conv2d_input in fp64.  Odd
    extents, groups, stride larger than the filter (positions that receive nothing), padding, a ReLU mask."""
    from cxxnet_b200 import ops
    L = ops.lib()
    torch.manual_seed(C * 7 + k * 3 + s)
    N = 3
    icg, ocg = C // G, OC // G
    OH, OW = (H + 2 * pad - k) // s + 1, (W + 2 * pad - k) // s + 1
    w = _bf(torch.randn(OC, icg, k, k, device="cuda") * (1.0 / (icg * k * k) ** 0.5))
    dy = _bf(torch.randn(N, OH, OW, OC, device="cuda"))
    kg = k * k * icg
    kpad = ops.round_up(kg, 8)
    wsh = torch.zeros(G * ocg, kpad, dtype=torch.bfloat16, device="cuda")
    wsh[:, :kg] = w.view(G, ocg, icg, k, k).permute(0, 1, 3, 4, 2).reshape(G * ocg, kg)
    assert L.cxn_conv_dgrad_strided_ok(k, k, s, pad, pad) == 1
    scratch = torch.zeros(int(L.cxn_conv_dgrad_strided_scratch(G, ocg, icg, k, k, s)) + 8, dtype=torch.bfloat16, device="cuda")
    dxr = torch.nn.grad.conv2d_input((N, C, H, W), w.double(), dy.double().permute(0, 3, 1, 2), stride=s, padding=pad,
                                     groups=G).float()
    for use_mask in (False, True):
        dx = torch.full((N * H * W, C), 3.0, dtype=torch.bfloat16, device="cuda")     # stale values must all be overwritten
        mask = _bf(torch.randn(N * H * W, C, device="cuda")) if use_mask else None
        ops.check(L.cxn_conv_dgrad_strided(dy.data_ptr(), N, OH, OW, OC, ocg, G, wsh.data_ptr(), kpad, icg, k, k, s, pad, pad,
                                           H, W, dx.data_ptr(), C, mask.data_ptr() if use_mask else None,
                                           scratch.data_ptr(), ops.stream()), "dgrad_strided")
        torch.cuda.synchronize()
        want = dxr
        if use_mask:
            want = dxr * (mask.float().view(N, H, W, C).permute(0, 3, 1, 2) > 0)
        got = dx.float().view(N, H, W, C).permute(0, 3, 1, 2)
        err = float((got - want).abs().max())
        assert err <= TOL_BF16_OUT * max(1.0, float(dxr.abs().max())), "strided dgrad err %g (mask %s)" % (err, use_mask)
