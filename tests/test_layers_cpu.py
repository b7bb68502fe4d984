"""Finite-difference checks of every layer's reference-backend math (float64), plus the shape
rules and quirks listed in SURVEY.md section 7.3 item 6."""
import math

import numpy as np
import pytest
import torch

from cxxnet_b200 import netconfig as nc
from cxxnet_b200.layers import LabelInfo, Node, create_layer
from cxxnet_b200.layers.base import pool_out_size


def build(type_name, params, in_shapes, n_out=1, self_loop=False, label=None):
    layer = create_layer(nc.get_layer_type(type_name))
    layer.rng = torch.Generator().manual_seed(7)
    info = LabelInfo({"label": 0})
    if label is not None:
        info.fields = [label]
    layer.label_info = info
    for k, v in params.items():
        layer.set_param(k, str(v))
    nodes_in = [Node("in%d" % i, i) for i in range(len(in_shapes))]
    for n, s in zip(nodes_in, in_shapes):
        n.shape = list(s)
    nodes_out = nodes_in if self_loop else [Node("out%d" % i, 10 + i) for i in range(n_out)]
    layer.init_connection(nodes_in, nodes_out)
    layer.init_model()
    for tag, w, g in layer.weights():
        layer.rebind(tag, w.double(), g.double())
    if getattr(layer, "wmat", None) is not None and not layer.weights():
        layer.wmat = layer.wmat.double()
    return layer, nodes_in, nodes_out


def run_fwd(layer, nodes_in, nodes_out, xs, is_train=True):
    layer.rng.manual_seed(123)
    for n, x in zip(nodes_in, xs):
        n.data = x.clone()
    layer.forward(is_train, nodes_in, nodes_out)
    return [n.data.clone() for n in nodes_out]


def fd_check(type_name, params, in_shapes, n_out=1, self_loop=False, label=None, tol=1e-5, nprobe=6, is_train=True,
             check_weights=True):
    torch.manual_seed(0)
    layer, nin, nout = build(type_name, params, in_shapes, n_out, self_loop, label)
    xs = [torch.randn(s, dtype=torch.float64) for s in in_shapes]
    outs = run_fwd(layer, nin, nout, xs, is_train)
    Gs = [torch.randn_like(o) for o in outs]

    def loss():
        o = run_fwd(layer, nin, nout, xs, is_train)
        return sum(float((a * g).sum()) for a, g in zip(o, Gs))
    # analytic
    run_fwd(layer, nin, nout, xs, is_train)
    for _, _, g in layer.weights():
        g.zero_()
    if self_loop:
        nin[0].data = Gs[0].clone() if type_name not in ("softmax", "l2_loss", "multi_logistic") else nin[0].data
    else:
        for n, g in zip(nout, Gs):
            n.data = g.clone()
    layer.backprop(True, nin, nout)
    gin = [n.data.clone() for n in nin]
    gws = [(tag, w, g.clone()) for tag, w, g in layer.weights()]
    eps = 1e-6
    rs = np.random.RandomState(1)
    for x, gx in zip(xs, gin):
        flat = x.view(-1)
        for idx in rs.choice(flat.numel(), size=min(nprobe, flat.numel()), replace=False):
            old = float(flat[idx])
            flat[idx] = old + eps; lp = loss()
            flat[idx] = old - eps; lm = loss()
            flat[idx] = old
            num = (lp - lm) / (2 * eps)
            assert abs(num - float(gx.view(-1)[idx])) <= tol * max(1.0, abs(num)), \
                "%s: d/dx mismatch %g vs %g" % (type_name, num, float(gx.view(-1)[idx]))
    for tag, w, gw in (gws if check_weights else []):
        flat = w.view(-1)
        for idx in rs.choice(flat.numel(), size=min(nprobe, flat.numel()), replace=False):
            old = float(flat[idx])
            flat[idx] = old + eps; lp = loss()
            flat[idx] = old - eps; lm = loss()
            flat[idx] = old
            num = (lp - lm) / (2 * eps)
            assert abs(num - float(gw.view(-1)[idx])) <= tol * max(1.0, abs(num)), \
                "%s: d/d%s mismatch %g vs %g" % (type_name, tag, num, float(gw.view(-1)[idx]))
    return layer


def test_fullc():
    fd_check("fullc", {"nhidden": 7, "init_sigma": 0.5}, [(4, 1, 1, 9)])


def test_conv_grouped_strided_padded():
    fd_check("conv", {"nchannel": 6, "ngroup": 2, "kernel_size": 3, "stride": 2, "pad": 1, "init_sigma": 0.5},
             [(2, 4, 7, 8)])
    fd_check("conv", {"nchannel": 4, "kernel_height": 2, "kernel_width": 3, "pad_y": 0, "pad_x": 1,
                      "random_type": "xavier"}, [(2, 3, 5, 6)])


@pytest.mark.parametrize("kind", ["max_pooling", "sum_pooling", "avg_pooling"])
def test_pooling_ceil_mode_clipped(kind):
    fd_check(kind, {"kernel_size": 3, "stride": 2}, [(2, 3, 8, 7)])


def test_pool_shape_rule_and_avg_divides_by_full_window():
    assert pool_out_size(55, 3, 2) == 27 and pool_out_size(13, 3, 2) == 6 and pool_out_size(8, 3, 2) == 4
    assert pool_out_size(5, 5, 5) == 1 and pool_out_size(6, 2, 4) == 2
    layer, nin, nout = build("avg_pooling", {"kernel_size": 3, "stride": 2}, [(1, 1, 4, 4)])
    (y,) = run_fwd(layer, nin, nout, [torch.ones(1, 1, 4, 4, dtype=torch.float64)])
    assert y.shape == (1, 1, 2, 2)
    assert y[0, 0, 0, 0] == pytest.approx(1.0) and y[0, 0, 1, 1] == pytest.approx(4.0 / 9.0)  # clipped 2x2 window / 9


def test_max_unpool_sends_gradient_to_every_tie():
    layer, nin, nout = build("max_pooling", {"kernel_size": 2, "stride": 2}, [(1, 1, 2, 2)])
    x = torch.tensor([[[[1.0, 1.0], [0.0, 1.0]]]], dtype=torch.float64)
    run_fwd(layer, nin, nout, [x])
    nout[0].data = torch.full((1, 1, 1, 1), 2.0, dtype=torch.float64)
    layer.backprop(True, nin, nout)
    assert nin[0].data.tolist() == [[[[2.0, 2.0], [0.0, 2.0]]]]


def test_relu_max_pooling():
    fd_check("relu_max_pooling", {"kernel_size": 2, "stride": 2}, [(2, 2, 6, 6)])


@pytest.mark.parametrize("kind,params", [("relu", {}), ("sigmoid", {}), ("tanh", {}), ("xelu", {"b": 3})])
def test_activations_in_place_semantics(kind, params):
    fd_check(kind, params, [(3, 2, 4, 4)])
    layer, nin, nout = build(kind, params, [(2, 1, 1, 5)])
    x = torch.randn(2, 1, 1, 5, dtype=torch.float64)
    (y,) = run_fwd(layer, nin, nout, [x])
    assert torch.equal(nin[0].data, y)       # the input node holds the activated value too


def test_lrn():
    fd_check("lrn", {"local_size": 5, "alpha": 0.1, "beta": 0.75, "knorm": 1}, [(2, 7, 3, 3)])


def test_dropout_scaling_and_eval_identity():
    layer, nin, nout = build("dropout", {"threshold": 0.5}, [(4, 1, 1, 1000)], self_loop=True)
    x = torch.ones(4, 1, 1, 1000, dtype=torch.float64)
    (y,) = run_fwd(layer, nin, nout, [x], is_train=True)
    vals = set(np.round(y.unique().tolist(), 6))
    assert vals == {0.0, 2.0}                                # mask / pkeep
    assert 0.4 < float((y > 0).double().mean()) < 0.6
    (y2,) = run_fwd(layer, nin, nout, [x], is_train=False)
    assert torch.equal(y2, x)
    fd_check("dropout", {"threshold": 0.3}, [(3, 1, 1, 20)], self_loop=True)


def test_flatten_split_concat():
    fd_check("flatten", {}, [(2, 3, 4, 5)])
    fd_check("split", {}, [(2, 3, 2, 2)], n_out=3)
    fd_check("ch_concat", {}, [(2, 3, 2, 2), (2, 1, 2, 2), (2, 2, 2, 2)])
    fd_check("concat", {}, [(2, 1, 1, 4), (2, 1, 1, 3)])


def test_bias_prelu_batchnorm_insanity():
    fd_check("bias", {"init_bias": 0.3}, [(3, 1, 1, 6)], self_loop=True)
    fd_check("prelu", {"init_slope": 0.25}, [(2, 3, 4, 4)])
    # with training noise the reference's slope gradient ignores the noise factor
    # (prelu_layer-inl.hpp:146,151), so only the input gradient is exact
    fd_check("prelu", {"init_slope": 0.3, "random": 0.2}, [(3, 1, 1, 8)], check_weights=False)
    fd_check("batch_norm", {"eps": 1e-3, "init_slope": 1.3, "init_bias": 0.2}, [(4, 3, 3, 3)], tol=1e-4)
    fd_check("batch_norm", {"eps": 1e-3}, [(6, 1, 1, 5)], tol=1e-4)
    fd_check("insanity", {"lb": 3, "ub": 8}, [(2, 2, 3, 3)])


def test_insanity_pooling_semantics():
    # keep = 1: identical to max pooling (values and gradient); eval mode is always plain max pooling.
    # (for keep < 1 the reference routes the gradient to the pooled position, not to the displaced
    # source pixel - insanity_pooling_layer-inl.hpp:162-214 - so it is not a true derivative)
    fd_check("insanity_max_pooling", {"kernel_size": 2, "stride": 2, "keep": 1.0}, [(2, 2, 6, 6)])
    layer, nin, nout = build("insanity_max_pooling", {"kernel_size": 3, "stride": 2, "keep": 0.5}, [(2, 3, 7, 7)])
    x = torch.randn(2, 3, 7, 7, dtype=torch.float64)
    (ye,) = run_fwd(layer, nin, nout, [x], is_train=False)
    ref, rin, rout = build("max_pooling", {"kernel_size": 3, "stride": 2}, [(2, 3, 7, 7)])
    (yr,) = run_fwd(ref, rin, rout, [x])
    assert torch.equal(ye, yr)
    (yt,) = run_fwd(layer, nin, nout, [x], is_train=True)
    assert yt.shape == yr.shape and not torch.equal(yt, yr)
    nout[0].data = torch.ones_like(yt)
    layer.backprop(True, nin, nout)
    assert nin[0].data.shape == x.shape and float(nin[0].data.sum()) >= yt.numel()


def test_batchnorm_uses_batch_statistics_at_test_time():
    layer, nin, nout = build("batch_norm", {"eps": 1e-5}, [(8, 2, 3, 3)])
    x = torch.randn(8, 2, 3, 3, dtype=torch.float64) * 3 + 5
    (y,) = run_fwd(layer, nin, nout, [x], is_train=False)
    assert abs(float(y.mean())) < 1e-6 and abs(float(y.var(unbiased=False)) - 1.0) < 1e-3


def test_fixconn(tmp_path):
    f = tmp_path / "w.txt"
    f.write_text("3 4 5\n0 0 1.0\n0 3 -2.0\n1 1 0.5\n2 2 3.0\n2 0 1.5\n")
    fd_check("fixconn", {"nhidden": 3, "fixconn_weight": str(f)}, [(2, 1, 1, 4)])


def test_loss_layers_scale_by_global_batch():
    label = torch.tensor([[1.0], [0.0], [2.0]])
    layer, nin, _ = build("softmax", {"batch_size": 12, "update_period": 2, "grad_scale": 3}, [(3, 1, 1, 4)],
                          self_loop=True, label=label)
    x = torch.randn(3, 1, 1, 4, dtype=torch.float64)
    nin[0].data = x.clone()
    layer.forward(True, nin, nin)
    p = torch.softmax(x.view(3, 4), 1)
    assert torch.allclose(nin[0].data.view(3, 4), p)
    layer.backprop(True, nin, nin)
    exp = p.clone()
    exp[torch.arange(3), label[:, 0].long()] -= 1.0
    assert torch.allclose(nin[0].data.view(3, 4), exp * (3.0 / (12 * 2)))
    lab2 = torch.rand(3, 4)
    for kind, tr in (("l2_loss", lambda v: v), ("multi_logistic", torch.sigmoid)):
        layer, nin, _ = build(kind, {"batch_size": 3}, [(3, 1, 1, 4)], self_loop=True, label=lab2)
        nin[0].data = x.clone()
        layer.forward(True, nin, nin)
        layer.backprop(True, nin, nin)
        assert torch.allclose(nin[0].data.view(3, 4), (tr(x.view(3, 4)) - lab2.double()) / 3.0)


def test_unimplemented_types_rejected_like_reference():
    for name in ("softplus", "maxout"):
        with pytest.raises(ValueError):
            create_layer(nc.get_layer_type(name))


def test_weight_init_modes():
    for rt, chk in (("gaussian", lambda w: abs(float(w.std()) - 0.01) < 0.003),
                    ("xavier", lambda w: float(w.abs().max()) <= math.sqrt(3.0 / (50 + 40)) + 1e-6),
                    ("kaiming", lambda w: abs(float(w.std()) - math.sqrt(2.0 / 40)) < 0.05)):
        layer = create_layer(nc.kFullConnect)
        layer.rng = torch.Generator().manual_seed(3)
        layer.set_param("nhidden", "40"); layer.set_param("random_type", rt)
        a, b = Node("a", 0), Node("b", 1)
        a.shape = [2, 1, 1, 50]
        layer.init_connection([a], [b])
        layer.init_model()
        assert layer.wmat.shape == (40, 50) and chk(layer.wmat), rt


@pytest.mark.parametrize("mode,kw,kw_eff", [("plain", 11, 11), ("pair", 11, 12), ("span", 11, 16), ("pair", 4, 4), ("span", 3, 16)])
def test_native_conv_first_layer_weight_layouts_round_trip(mode, kw, kw_eff):
    """internal weight layouts of the native first convolution (channel pitch 4; pixel pairs; 16-pixel
    spans): checkpoint layout -> internal -> checkpoint is the identity and the added filter columns /
    the fourth channel are structural zeros"""
    from cxxnet_b200.layers import native
    l = native.ConvolutionLayer()
    for k, v in (("nchannel", "24"), ("kernel_size", str(kw)), ("stride", "2")):
        l.set_param(k, v)
    l.param.num_input_channel = 3
    l.cin_pad = 4
    l.pair = mode in ("pair", "span")
    l.span = mode == "span"
    torch.manual_seed(3)
    w = torch.randn(24, 3 * kw * kw)
    wi = l.weight_to_internal("wmat", w)
    assert wi.shape == (1, 24, kw * kw_eff * 4)
    w5 = wi.reshape(24, kw, kw_eff, 4)
    assert float(w5[:, :, kw:, :].abs().sum()) == 0.0 and float(w5[..., 3].abs().sum()) == 0.0
    # (oc, c, kh, kw) of the checkpoint is (oc, kh, kw, c) inside
    assert torch.equal(w5[:, :, :kw, :3], w.reshape(24, 3, kw, kw).permute(0, 2, 3, 1))
    back = l.weight_from_internal("wmat", wi)
    assert torch.equal(back.reshape(24, -1), w)
