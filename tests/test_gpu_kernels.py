"""Stand-alone sm_100a kernels (pooling, LRN, activations, dropout, reductions, loss, updater,
layout) against plain PyTorch fp32 references on identical bf16-rounded inputs."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def bf(t):
    return t.to(torch.bfloat16)


def nhwc(t):      # (n,c,h,w) fp32 -> bf16 [n*h*w, c]
    n, c, h, w = t.shape
    return bf(t.permute(0, 2, 3, 1).reshape(n * h * w, c)).contiguous()


def from_nhwc(m, n, c, h, w):
    return m.float().view(n, h, w, c).permute(0, 3, 1, 2)


@pytest.mark.parametrize("mode,name", [(0, "max"), (1, "sum"), (2, "avg")])
@pytest.mark.parametrize("shape,k,s", [((3, 16, 13, 13), 3, 2), ((2, 5, 9, 8), 2, 2), ((2, 96, 55, 55), 3, 2)])
def test_pool_fwd_bwd(mode, name, shape, k, s):
    from cxxnet_b200 import ops
    from cxxnet_b200.layers.zoo import pool_backward, pool_forward
    L = ops.lib()
    torch.manual_seed(1)
    n, c, h, w = shape
    x = bf(torch.randn(shape, device="cuda")).float()
    xr = nhwc(x)
    yref = pool_forward(x, k, k, s, name)
    oh, ow = yref.shape[2], yref.shape[3]
    y = torch.zeros(n * oh * ow, c, dtype=torch.bfloat16, device="cuda")
    ops.check(L.cxn_pool_fwd(xr.data_ptr(), y.data_ptr(), n, h, w, c, k, k, s, oh, ow, mode, 0, ops.stream()), "pool")
    got = from_nhwc(y, n, c, oh, ow)
    assert torch.allclose(got, bf(yref).float(), atol=1e-2, rtol=1e-2)
    g = bf(torch.randn_like(yref)).float()
    dxref = pool_backward(x, bf(yref).float() if mode == 0 else yref, g, k, k, s, name)
    dx = torch.zeros_like(xr)
    ops.check(L.cxn_pool_bwd(xr.data_ptr(), y.data_ptr(), nhwc(g).data_ptr(), dx.data_ptr(), n, h, w, c, k, k, s, oh,
                             ow, mode, 0, ops.stream()), "pool_bwd")
    gotdx = from_nhwc(dx, n, c, h, w)
    assert torch.allclose(gotdx, dxref, atol=2e-2, rtol=2e-2)


@pytest.mark.parametrize("shape,k,s", [((6, 96, 55, 55), 3, 2), ((5, 256, 27, 27), 3, 2), ((4, 256, 13, 13), 3, 2),
                                       ((3, 16, 9, 8), 2, 2)])
def test_pool_bwd_with_fused_bias_gradient(shape, k, s):
    """cxn_pool_bwd_bias: the same dx as cxn_pool_bwd (ReLU-masked max pooling) and, in the same pass, the column
    sums of dx = the bias gradient of the convolution in front (plain fp32 PyTorch sum of the stored bf16 values);
    a second call accumulates onto the first"""
    from cxxnet_b200 import ops
    from cxxnet_b200.layers.zoo import pool_forward
    L = ops.lib()
    torch.manual_seed(7)
    n, c, h, w = shape
    x = bf(torch.randn(shape, device="cuda")).float()
    xr = nhwc(x)
    yref = pool_forward(torch.relu(x), k, k, s, "max")
    oh, ow = yref.shape[2:]
    y = torch.zeros(n * oh * ow, c, dtype=torch.bfloat16, device="cuda")
    ops.check(L.cxn_pool_fwd(xr.data_ptr(), y.data_ptr(), n, h, w, c, k, k, s, oh, ow, 0, 1, ops.stream()), "pool")
    g = nhwc(bf(torch.randn_like(yref)).float())
    dx0, dx1 = torch.zeros_like(xr), torch.zeros_like(xr)
    ops.check(L.cxn_pool_bwd(xr.data_ptr(), y.data_ptr(), g.data_ptr(), dx0.data_ptr(), n, h, w, c, k, k, s, oh, ow, 0, 1,
                             ops.stream()), "pool_bwd")
    gb = torch.zeros(c, device="cuda")
    for rep in range(2):
        ops.check(L.cxn_pool_bwd_bias(xr.data_ptr(), y.data_ptr(), g.data_ptr(), dx1.data_ptr(), gb.data_ptr(), n, h, w, c, k,
                                      k, s, oh, ow, 0, 1, ops.stream()), "pool_bwd_bias")
        torch.cuda.synchronize()
        assert torch.equal(dx0, dx1)
        want = (rep + 1) * dx0.float().sum(dim=0)
        assert torch.allclose(gb, want, atol=1e-3 * max(1.0, float(want.abs().max())), rtol=1e-4), (rep, float((gb - want).abs().max()))


@pytest.mark.parametrize("relu", [0, 1])
@pytest.mark.parametrize("shape,k", [((2, 16, 10, 12), 3), ((2, 8, 11, 9), 4), ((3, 32, 28, 28), 2), ((2, 96, 55, 55), 3),
                                     ((2, 256, 13, 13), 3), ((1, 8, 3, 3), 3), ((2, 8, 5, 4), 1)])
def test_pool_bwd_stride2_blocks_with_ties(shape, k, relu):
    """stride-2 max-pool backward (one thread per 2 x 2 input block, pool_bwd_quad_kernel): inputs drawn from a few
    values, so windows hold ties and EVERY maximal input must receive the gradient; odd and even extents, window
    sizes 1-4, with and without the fused ReLU mask"""
    from cxxnet_b200 import ops
    from cxxnet_b200.layers.zoo import pool_backward, pool_forward
    L = ops.lib()
    torch.manual_seed(k * 10 + relu)
    n, c, h, w = shape
    x = (torch.randint(-2, 3, shape, device="cuda").float() * 0.5)
    xr = nhwc(x)
    xin = torch.relu(x) if relu else x
    yref = pool_forward(xin, k, k, 2, "max")
    oh, ow = yref.shape[2:]
    y = torch.zeros(n * oh * ow, c, dtype=torch.bfloat16, device="cuda")
    ops.check(L.cxn_pool_fwd(xr.data_ptr(), y.data_ptr(), n, h, w, c, k, k, 2, oh, ow, 0, relu, ops.stream()), "pool")
    assert torch.equal(from_nhwc(y, n, c, oh, ow), yref)
    g = bf(torch.randint(-4, 5, yref.shape, device="cuda").float() * 0.25).float()       # sums stay exact in bf16
    dxref = pool_backward(xin, yref, g, k, k, 2, "max")
    if relu:
        dxref = dxref * (x > 0)
    dx = torch.full_like(xr, 7.0)
    ops.check(L.cxn_pool_bwd(xr.data_ptr(), y.data_ptr(), nhwc(g).data_ptr(), dx.data_ptr(), n, h, w, c, k, k, 2, oh,
                             ow, 0, relu, ops.stream()), "pool_bwd")
    assert torch.equal(from_nhwc(dx, n, c, h, w), dxref)


def test_pool_relu_mask_variants():
    from cxxnet_b200 import ops
    from cxxnet_b200.layers.zoo import pool_backward, pool_forward
    L = ops.lib()
    n, c, h, w, k, s = 2, 8, 9, 9, 3, 2
    x = bf(torch.randn(n, c, h, w, device="cuda")).float()
    xr = nhwc(x)
    yref = pool_forward(torch.relu(x), k, k, s, "max")
    oh, ow = yref.shape[2:]
    y = torch.zeros(n * oh * ow, c, dtype=torch.bfloat16, device="cuda")
    ops.check(L.cxn_pool_fwd(xr.data_ptr(), y.data_ptr(), n, h, w, c, k, k, s, oh, ow, 0, 1, ops.stream()), "pool")
    assert torch.allclose(from_nhwc(y, n, c, oh, ow), yref, atol=1e-2)
    g = bf(torch.randn_like(yref)).float()
    dxref = pool_backward(torch.relu(x), yref, g, k, k, s, "max") * (x > 0)
    dx = torch.zeros_like(xr)
    ops.check(L.cxn_pool_bwd(xr.data_ptr(), y.data_ptr(), nhwc(g).data_ptr(), dx.data_ptr(), n, h, w, c, k, k, s, oh,
                             ow, 0, 1, ops.stream()), "pool_bwd")
    assert torch.allclose(from_nhwc(dx, n, c, h, w), dxref, atol=2e-2, rtol=2e-2)


@pytest.mark.parametrize("shape,nsize", [((4, 96, 27, 27), 5), ((3, 256, 13, 13), 5), ((2, 24, 5, 5), 3),
                                         ((2, 20, 6, 6), 5), ((2, 7, 4, 4), 3)])
def test_lrn_fwd_bwd(shape, nsize):
    from cxxnet_b200 import ops
    from cxxnet_b200.layers.zoo import _chpool_sum
    L = ops.lib()
    torch.manual_seed(2)
    n, c, h, w = shape
    alpha, beta, knorm = 0.001 * 50, 0.75, 1.0
    x = bf(torch.randn(shape, device="cuda") * 3).float()
    g = bf(torch.randn(shape, device="cuda")).float()
    norm = _chpool_sum(x * x, nsize) * (alpha / nsize) + knorm
    yref = x * norm.pow(-beta)
    dxref = g * norm.pow(-beta) - 2 * beta * (alpha / nsize) * _chpool_sum(g * x * norm.pow(-beta - 1), nsize) * x
    xr, gr = nhwc(x), nhwc(g)
    y = torch.zeros_like(xr)
    dx = torch.zeros_like(xr)
    ops.check(L.cxn_lrn_fwd(xr.data_ptr(), y.data_ptr(), n * h * w, c, nsize, alpha, beta, knorm, ops.stream()), "lrn")
    ops.check(L.cxn_lrn_bwd(xr.data_ptr(), gr.data_ptr(), dx.data_ptr(), n * h * w, c, nsize, alpha, beta, knorm,
                            ops.stream()), "lrn_bwd")
    assert torch.allclose(from_nhwc(y, n, c, h, w), yref, atol=2e-2, rtol=2e-2)
    assert torch.allclose(from_nhwc(dx, n, c, h, w), dxref, atol=2e-2, rtol=2e-2)


@pytest.mark.parametrize("op,fn,gfn", [
    (0, torch.relu, lambda y: (y > 0).float()),
    (1, torch.sigmoid, lambda y: y * (1 - y)),
    (2, torch.tanh, lambda y: 1 - y * y),
    (3, lambda x: torch.where(x > 0, x, x / 4.0), lambda y: torch.where(y > 0, torch.ones_like(y), torch.full_like(y, 0.25)))])
def test_activations(op, fn, gfn):
    from cxxnet_b200 import ops
    L = ops.lib()
    rows, cols, ld = 37, 50, 56
    x = torch.zeros(rows, ld, dtype=torch.bfloat16, device="cuda")
    x[:, :cols] = bf(torch.randn(rows, cols, device="cuda"))
    y = torch.zeros_like(x)
    ops.check(L.cxn_act_fwd(x.data_ptr(), y.data_ptr(), rows, cols, ld, op, 4.0, ops.stream()), "act")
    ref = fn(x[:, :cols].float())
    assert torch.allclose(y[:, :cols].float(), ref, atol=1e-2, rtol=1e-2)
    assert float(y[:, cols:].abs().max()) == 0.0
    g = torch.zeros_like(x)
    g[:, :cols] = bf(torch.randn(rows, cols, device="cuda"))
    dx = torch.zeros_like(x)
    ops.check(L.cxn_act_bwd(y.data_ptr(), g.data_ptr(), dx.data_ptr(), rows, cols, ld, op, 4.0, ops.stream()), "act_bwd")
    assert torch.allclose(dx[:, :cols].float(), gfn(y[:, :cols].float()) * g[:, :cols].float(), atol=1e-2, rtol=2e-2)


def test_dropout_mask_is_reproducible_and_scaled():
    from cxxnet_b200 import ops
    L = ops.lib()
    rows, cols = 64, 4096
    x = torch.ones(rows, cols, dtype=torch.bfloat16, device="cuda")
    cnt = torch.zeros(1, dtype=torch.int64, device="cuda")
    y1, y2, y3 = torch.zeros_like(x), torch.zeros_like(x), torch.zeros_like(x)
    ops.check(L.cxn_dropout(x.data_ptr(), y1.data_ptr(), rows, cols, cols, 0.5, 77, cnt.data_ptr(), ops.stream()), "d")
    ops.check(L.cxn_dropout(x.data_ptr(), y2.data_ptr(), rows, cols, cols, 0.5, 77, cnt.data_ptr(), ops.stream()), "d")
    ops.check(L.cxn_counter_add(cnt.data_ptr(), 1, ops.stream()), "c")
    ops.check(L.cxn_dropout(x.data_ptr(), y3.data_ptr(), rows, cols, cols, 0.5, 77, cnt.data_ptr(), ops.stream()), "d")
    assert torch.equal(y1, y2) and not torch.equal(y1, y3)              # same (seed, step) -> same mask
    assert set(y1.float().unique().tolist()) == {0.0, 2.0}
    assert abs(float((y1 > 0).float().mean()) - 0.5) < 0.01


@pytest.mark.parametrize("rows,cols", [(100, 10), (30000, 96), (5000, 256), (777, 1000), (256, 4096), (200003, 96),
                                       (43264, 384)])
def test_colsum(rows, cols):
    from cxxnet_b200 import ops
    L = ops.lib()
    ld = ops.round_up(cols, 8)
    x = torch.zeros(rows, ld, dtype=torch.bfloat16, device="cuda")
    x[:, :cols] = bf(torch.randn(rows, cols, device="cuda"))
    out = torch.full((cols,), 5.0, device="cuda")
    ops.check(L.cxn_colsum_bf16(x.data_ptr(), rows, cols, ld, out.data_ptr(), 0, ops.stream()), "colsum")
    ref = x[:, :cols].float().sum(0)
    assert torch.allclose(out, ref, atol=1e-2 * rows ** 0.5, rtol=1e-3)
    ops.check(L.cxn_colsum_bf16(x.data_ptr(), rows, cols, ld, out.data_ptr(), 1, ops.stream()), "colsum")
    assert torch.allclose(out, 2 * ref, atol=2e-2 * rows ** 0.5, rtol=1e-3)


def test_softmax_and_loss_gradients():
    from cxxnet_b200 import ops
    L = ops.lib()
    b, n, ld = 33, 1000, 1000
    x = torch.randn(b, ld, device="cuda") * 3
    p = x.clone()
    ops.check(L.cxn_softmax_rows(p.data_ptr(), b, n, ld, ops.stream()), "softmax")
    ref = torch.softmax(x, 1)
    assert torch.allclose(p, ref, atol=1e-5)
    label = torch.randint(0, n, (b, 1), device="cuda").float()
    g = torch.zeros(b, ld, dtype=torch.bfloat16, device="cuda")
    ops.check(L.cxn_loss_grad(p.data_ptr(), ld, label.data_ptr(), 1, g.data_ptr(), ld, b, n, 0.25, 0, ops.stream()), "g")
    gr = ref.clone()
    gr[torch.arange(b), label[:, 0].long()] -= 1
    assert torch.allclose(g.float(), gr * 0.25, atol=2e-3, rtol=1e-2)
    dense = torch.rand(b, 7, device="cuda")
    q = torch.randn(b, 8, device="cuda")
    s = q.clone()
    ops.check(L.cxn_sigmoid_rows(s.data_ptr(), b, 7, 8, ops.stream()), "sig")
    assert torch.allclose(s[:, :7], torch.sigmoid(q[:, :7]), atol=1e-5)
    g2 = torch.zeros(b, 8, dtype=torch.bfloat16, device="cuda")
    ops.check(L.cxn_loss_grad(s.data_ptr(), 8, dense.data_ptr(), 7, g2.data_ptr(), 8, b, 7, 1.0, 1, ops.stream()), "g")
    assert torch.allclose(g2[:, :7].float(), s[:, :7] - dense, atol=1e-2)


def test_layout_converters_roundtrip():
    from cxxnet_b200 import ops
    L = ops.lib()
    for (n, c, h, w, ldc) in [(3, 3, 17, 19, 4), (2, 96, 7, 7, 96), (2, 5, 4, 6, 5), (2, 1, 28, 28, 4)]:
        x = torch.randn(n, c, h, w, device="cuda")
        d = torch.zeros(n * h * w, ldc, dtype=torch.bfloat16, device="cuda")
        ops.check(L.cxn_nchw_f32_to_nhwc_bf16(x.data_ptr(), d.data_ptr(), n, c, h * w, ldc, ops.stream()), "to")
        assert torch.equal(d[:, :c].float().view(n, h, w, c).permute(0, 3, 1, 2), bf(x).float())
        if ldc > c:
            assert float(d[:, c:].abs().max()) == 0.0
        back = torch.zeros_like(x)
        ops.check(L.cxn_nhwc_bf16_to_nchw_f32(d.data_ptr(), back.data_ptr(), n, c, h * w, ldc, ops.stream()), "from")
        assert torch.equal(back, bf(x).float())
    # flatten (c,h,w) order and back
    n, c, hw = 4, 24, 36
    src = bf(torch.randn(n, hw, c, device="cuda")).contiguous()
    ld = ops.round_up(c * hw, 8)
    dst = torch.zeros(n, ld, dtype=torch.bfloat16, device="cuda")
    ops.check(L.cxn_bf16_hwc_chw(src.data_ptr(), dst.data_ptr(), n, c, hw, hw * c, ld, c, hw, 1, ops.stream()), "f")
    assert torch.equal(dst[:, :c * hw].view(n, c, hw), src.permute(0, 2, 1))
    back = torch.zeros_like(src)
    ops.check(L.cxn_bf16_hwc_chw(dst.data_ptr(), back.data_ptr(), n, c, hw, ld, hw * c, c, hw, 0, ops.stream()), "u")
    assert torch.equal(back, src)


@pytest.mark.parametrize("kind", ["sgd", "nag", "adam"])
def test_arena_update_kernel_matches_reference_updaters(kind):
    from cxxnet_b200.parallel.arena import ParamArena
    from cxxnet_b200.updaters import UpdaterParam, create_updater
    dev = torch.device("cuda")
    arena = ParamArena(dev, kind)
    shapes = [(37, 50, 56), (1, 300, 0), (4, 1024, 1024)]
    ups = []
    for i, (r, c, sld) in enumerate(shapes):
        up = UpdaterParam("wmat" if sld else "bias")
        for k, v in [("lr", "0.05"), ("wd", "0.01" if sld else "0"), ("momentum", "0.9"), ("clip_gradient", "0.5" if i == 2 else "0"),
                     ("lr:schedule", "expdecay"), ("lr:gamma", "0.5"), ("lr:step", "3")]:
            up.set_param(k, v)
        ups.append(up)
        arena.add(None, i, up.tag, r, c, sld, up)
    arena.finalize()
    torch.manual_seed(3)
    refs = []
    for s in arena.segments:
        w = torch.randn(s.n, device=dev)
        arena.w_view(s).copy_(w)
        wr, gr = w.clone(), torch.zeros(s.n, device=dev)
        u = create_updater(kind, wr, gr, s.tag)
        u.param.__dict__.update({k: v for k, v in s.param.__dict__.items()})
        u.param.silent = 1
        u.init()
        refs.append((u, wr, gr))
    for epoch in range(4):
        for s, (u, wr, gr) in zip(arena.segments, refs):
            g = torch.randn(s.n, device=dev)
            if epoch == 2 and kind == "sgd" and s.param.clip_gradient:
                g[0] = float("nan")          # only the sgd updater clips (and maps NaN to 0)
            arena.g_view(s).copy_(g)
            gr.copy_(g)
        arena.refresh_hyper(epoch)
        arena.update()
        for s, (u, wr, gr) in zip(arena.segments, refs):
            u.update(epoch)
            assert torch.allclose(arena.w_view(s), wr, atol=1e-5, rtol=1e-5), (kind, epoch, s.tag)
            assert float(arena.g_view(s).abs().max()) == 0.0
            if s.shadow_off >= 0:
                sh = arena.shadow_view(s).view(s.rows, s.shadow_ld)[:, :s.cols]
                assert torch.equal(sh, arena.w_view(s).view(s.rows, s.cols).to(torch.bfloat16))


@pytest.mark.parametrize("rows,cols", [(256, 1000), (37, 10), (64, 1), (5, 33)])
def test_metric_error_kernel_counts_first_maximum_mismatches(rows, cols):
    """the fused error metric against the reference rule (first maximum wins; width 1: score > 0 is class 1)"""
    g = torch.Generator().manual_seed(rows * 131 + cols)
    scores = torch.randn(rows, cols, generator=g)
    if cols > 4:
        scores[:, 3] = scores[:, 1]                      # exact ties: the first maximum must win
        scores[: rows // 2, 1] = 10.0
        scores[: rows // 2, 3] = 10.0
    nclass = max(cols, 2)
    label = torch.randint(0, nclass, (rows, 1), generator=g).float()
    ld = (cols + 3) // 4 * 4
    p = torch.zeros(rows, ld, device="cuda")
    p[:, :cols] = scores.cuda()
    lab = label.cuda()
    acc = torch.zeros((), dtype=torch.float64, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    from cxxnet_b200 import ops
    L = ops.lib()
    for _ in range(2):                                   # accumulates
        assert L.cxn_metric_error(p.data_ptr(), ld, lab.data_ptr(), 1, rows, cols, acc.data_ptr(), st) == 0
    torch.cuda.synchronize()
    if cols == 1:
        pred = (scores[:, 0] > 0).long()
    else:
        mx = scores.max(dim=1, keepdim=True).values
        pred = (scores == mx).float().argmax(dim=1)
    expect = 2.0 * float((pred != label[:, 0].long()).sum())
    assert float(acc.item()) == expect
