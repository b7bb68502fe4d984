"""Net-level differential tests: every native sm_100a layer against the reference-math backend
(the `pairtest` idea of reference src/layer/pairtest_layer-inl.hpp applied to whole nets)."""
import pytest
import torch

from util import compare_backends, make_batch

pytestmark = pytest.mark.gpu

HEAD = """
batch_size = %d
input_shape = %s
eta = 0.1
momentum = 0.9
random_type = xavier
metric = error
"""


def _conf(body, batch, shape):
    return "netconfig=start\n" + body + "\nnetconfig=end\n" + HEAD % (batch, shape)


def test_mlp_sigmoid_softmax():
    body = """
layer[+1] = fullc:fc1
  nhidden = 100
layer[+1] = sigmoid
layer[+1] = fullc:fc2
  nhidden = 10
layer[+0] = softmax
"""
    compare_backends(_conf(body, 100, "1,1,784"), make_batch((100, 1, 1, 784)))


def test_fullc_relu_fused_tanh_l2():
    body = """
layer[+1] = fullc:fc1
  nhidden = 264
layer[+1] = relu
layer[+1] = fullc:fc2
  nhidden = 72
layer[+1] = tanh
layer[+1] = fullc:fc3
  nhidden = 5
layer[+0] = l2_loss
  target = vec
"""
    compare_backends(_conf(body, 37, "1,1,300") + "label_vec[0,5) = vec\n", make_batch((37, 1, 1, 300), dense_label=5))


def test_conv_stride_pad_pool_flatten():
    body = """
layer[+1] = conv:cv1
  kernel_size = 3
  pad = 1
  stride = 2
  nchannel = 32
layer[+1] = max_pooling
  kernel_size = 3
  stride = 2
layer[+1] = flatten
layer[+1] = fullc:fc1
  nhidden = 10
layer[+0] = softmax
"""
    # bf16 activations create exact ties inside max-pool windows (every tied input receives the
    # gradient, as in the reference's unpool), so gradients below a max-pool carry more noise
    compare_backends(_conf(body, 20, "1,28,28"), make_batch((20, 1, 28, 28)), check_nodes=[1, 2, 4], tol_grad=0.12)


@pytest.mark.parametrize("span", [1, 0])
@pytest.mark.parametrize("k,stride,shape", [(11, 4, (3, 67, 67)), (4, 2, (3, 30, 31)), (3, 2, (1, 21, 20))])
def test_conv_first_layer_spans_and_pixel_pairs(k, stride, shape, span, monkeypatch):
    """first layer, <= 4 input channels, even stride, no padding.  span=1: 16-pixel x 4-channel runs of the
    pitch-4 image are 64-channel positions fetched by im2col-mode TMA; span=0: pixel pairs (16-byte gather
    granules).  Either way the unused filter columns are structural zeros of the internal weight layout."""
    monkeypatch.setenv("CXN_CONV_SPAN", str(span))
    monkeypatch.setenv("CXN_CONV_S2D", "0")        # (stride-4 layers would otherwise take the space-to-depth route)
    body = """
layer[+1] = conv:cv1
  kernel_size = %d
  stride = %d
  nchannel = 24
layer[+1] = relu
layer[+1] = flatten
layer[+1] = fullc:fc1
  nhidden = 10
layer[+0] = softmax
""" % (k, stride)
    shp = "%d,%d,%d" % shape
    errs, ref, b2 = compare_backends(_conf(body, 9, shp), make_batch((9,) + shape), return_trainers=True,
                                    tol_grad=0.1)   # bias gradients are sums of cancelling bf16 terms
    layer = b2.net.connections[0].layer
    assert layer.pair and layer.span == bool(span)
    kw_eff = 16 if layer.span else 2 * ((k + 1) // 2)
    if kw_eff > k:
        # the structural zeros must not collect gradient (their patch values are real pixels)
        g = layer.bind["wmat"].g.reshape(24, k, kw_eff, 4)
        assert float(g[:, :, k:, :].abs().max()) == 0.0
        assert float(g[:, :, :k, :shape[0]].abs().max()) > 0.0


def test_conv_groups_relu_lrn():
    body = """
layer[+1] = conv:cv1
  kernel_size = 5
  stride = 2
  nchannel = 16
layer[+1] = relu
layer[+1] = max_pooling
  kernel_size = 3
  stride = 2
layer[+1] = lrn
  local_size = 5
  alpha = 0.001
  beta = 0.75
  knorm = 1
layer[+1] = conv:cv2
  kernel_size = 3
  pad = 1
  ngroup = 2
  nchannel = 32
layer[+1] = relu
layer[+1] = conv:cv3
  kernel_size = 3
  pad = 1
  ngroup = 2
  nchannel = 48
layer[+1] = relu
layer[+1] = avg_pooling
  kernel_size = 2
  stride = 2
layer[+1] = flatten
layer[+1] = fullc:fc1
  nhidden = 7
layer[+0] = multi_logistic
  target = vec
"""
    compare_backends(_conf(body, 6, "3,37,37") + "label_vec[0,7) = vec\n", make_batch((6, 3, 37, 37), dense_label=7),
                     tol_grad=0.12)


def test_pool_variants_and_xelu():
    body = """
layer[+1] = conv:cv1
  kernel_size = 3
  nchannel = 8
layer[+1] = xelu
  b = 4
layer[+1] = sum_pooling
  kernel_size = 2
  stride = 2
layer[+1] = conv:cv2
  kernel_size = 2
  nchannel = 16
layer[+1] = relu_max_pooling
  kernel_size = 3
  stride = 2
layer[+1] = flatten
layer[+1] = fullc:fc1
  nhidden = 4
layer[+0] = softmax
"""
    compare_backends(_conf(body, 5, "2,19,19"), make_batch((5, 2, 19, 19), nclass=4), tol_grad=0.15)


def test_bn_prelu_split_ch_concat_on_images():
    body = """
layer[0->1] = conv:cv1
  kernel_size = 3
  pad = 1
  nchannel = 8
layer[1->2] = batch_norm:bn1
layer[2->3] = prelu:pr1
layer[3->4,5] = split
layer[4->6] = conv:cv2
  kernel_size = 3
  pad = 1
  nchannel = 8
layer[5->7] = conv:cv3
  kernel_size = 1
  nchannel = 8
layer[6,7->8] = ch_concat
layer[8->9] = flatten
layer[9->10] = fullc:fc1
  nhidden = 6
layer[10->10] = softmax
"""
    compare_backends(_conf(body, 4, "3,9,9") + "eps = 1e-5\n", make_batch((4, 3, 9, 9), nclass=6), tol_grad=8e-2,
                     skip=("grad[0:bias]",))  # analytically zero: batch_norm removes the conv bias


def test_bias_bn_prelu_split_concat_on_matrices():
    body = """
layer[0->1] = fullc:fc1
  nhidden = 32
  no_bias = 1
layer[1->1] = bias:b1
  init_bias = 0.3
layer[1->2] = batch_norm:bn1
layer[2->3] = prelu:pr1
layer[3->4,5] = split
layer[4->6] = fullc:fa
  nhidden = 16
layer[5->7] = fullc:fb
  nhidden = 24
layer[6,7->8] = concat
layer[8->9] = fullc:fc2
  nhidden = 6
layer[9->9] = softmax
"""
    errs, ref, b2 = compare_backends(_conf(body, 24, "1,1,40") + "eps = 1e-5\n", make_batch((24, 1, 1, 40), nclass=6),
                                     tol_grad=8e-2, skip=("grad[1:bias]",), return_trainers=True)  # bn removes the bias
    assert all(c.native for c in b2.net.connections), "every layer of this net has a native implementation"


def test_update_matches_reference_sgd_step():
    """weights after 3 sgd steps (momentum, wd, per-tag lr) agree with the reference backend."""
    from util import copy_weights, make_trainer, rel_err
    body = """
layer[+1] = fullc:fc1
  nhidden = 64
layer[+1] = relu
layer[+1] = fullc:fc2
  nhidden = 10
layer[+0] = softmax
"""
    conf = _conf(body, 32, "1,1,50") + "wmat:wd = 0.001\nbias:lr = 0.2\n"
    ref = make_trainer(conf, dev="gpu:0", backend="ref")
    b2 = make_trainer(conf, dev="gpu:0", backend="b200")
    copy_weights(ref, b2)
    for s in range(3):
        batch = make_batch((32, 1, 1, 50), seed=s)
        ref.update(batch)
        b2.update(batch)
    for name in ("fc1", "fc2"):
        for tag in ("wmat", "bias"):
            wr, _ = ref.get_weight(name, tag)
            wb, _ = b2.get_weight(name, tag)
            assert rel_err(torch.from_numpy(wr), torch.from_numpy(wb)) < 2e-2, (name, tag)


def test_insanity_and_fixconn_native(tmp_path):
    """insanity with lb == ub is a deterministic leaky ReLU (x / lb below zero), so the Philox stream of the
    native kernel and torch's generator in the reference math give the same answer; fixconn multiplies by a
    file-given sparse matrix and passes gradients through it"""
    import random
    rnd = random.Random(5)
    nout, nin = 24, 40
    entries = sorted({(rnd.randrange(nout), rnd.randrange(nin)) for _ in range(300)})
    wfile = tmp_path / "fix.txt"
    wfile.write_text("%d %d %d\n" % (nout, nin, len(entries)) +
                     "\n".join("%d %d %.4f" % (r, c, rnd.uniform(-1, 1)) for r, c in entries) + "\n")
    body = """
layer[0->1] = fullc:fc1
  nhidden = 40
layer[1->2] = insanity:in1
  lb = 4
  ub = 4
layer[2->3] = fixconn:fx1
  nhidden = 24
  fixconn_weight = %s
layer[3->4] = fullc:fc2
  nhidden = 6
layer[4->4] = softmax
""" % wfile
    errs, ref, b2 = compare_backends(_conf(body, 20, "1,1,30"), make_batch((20, 1, 1, 30), nclass=6), tol_grad=8e-2,
                                     return_trainers=True)
    assert all(c.native for c in b2.net.connections)


def test_insanity_pooling_native():
    """keep = 1 never displaces, so training mode must equal plain max pooling on both backends; the
    displacement kernel itself is checked against its definition: every output element is the input at the
    same position or at one of its four (clamped) neighbours, with roughly the configured frequencies"""
    body = """
layer[+1] = conv:cv1
  kernel_size = 3
  pad = 1
  nchannel = 16
layer[+1] = insanity_max_pooling
  kernel_size = 3
  stride = 2
  keep = 1.0
layer[+1] = flatten
layer[+1] = fullc:fc1
  nhidden = 10
layer[+0] = softmax
"""
    errs, ref, b2 = compare_backends(_conf(body, 8, "3,15,15"), make_batch((8, 3, 15, 15)), tol_grad=0.12,
                                     return_trainers=True)
    assert all(c.native for c in b2.net.connections)
    from cxxnet_b200 import ops
    L = ops.lib()
    N, H, W, C = 3, 9, 11, 16
    x = torch.arange(N * H * W * C, device="cuda", dtype=torch.float32).reshape(N, H, W, C)
    # values identify their source position: keep them exactly representable in bf16 by using small integers
    pos = (torch.arange(H, device="cuda").view(1, H, 1, 1) * 16 + torch.arange(W, device="cuda").view(1, 1, W, 1)).expand(N, H, W, C)
    xin = pos.to(torch.bfloat16).contiguous()
    out = torch.zeros_like(xin)
    counter = torch.zeros(1, dtype=torch.int64, device="cuda")
    assert L.cxn_insanity_displace(xin.data_ptr(), out.data_ptr(), N, H, W, C, C, 0.6, 12345, counter.data_ptr(),
                                   torch.cuda.current_stream().cuda_stream) == 0
    torch.cuda.synchronize()
    o = out.float().long()
    ys, xs = o // 16, o % 16
    y0 = torch.arange(H, device="cuda").view(1, H, 1, 1).expand(N, H, W, C)
    x0 = torch.arange(W, device="cuda").view(1, 1, W, 1).expand(N, H, W, C)
    dy, dx = ys - y0, xs - x0
    assert bool(((dy.abs() + dx.abs()) <= 1).all())                       # itself or a direct neighbour
    interior = (y0 > 0) & (y0 < H - 1) & (x0 > 0) & (x0 < W - 1)
    kept = float(((dy == 0) & (dx == 0) & interior).sum()) / float(interior.sum())
    assert 0.5 < kept < 0.7                                                 # keep = 0.6
    up = float(((dy == -1) & interior).sum()) / float(interior.sum())
    assert 0.06 < up < 0.14                                                 # (1 - keep) / 4 = 0.1


def test_shared_layers_stay_native():
    """`share[tag]` (reference neural_net-inl.hpp:238-244): a fully connected autoencoder with tied weights and a
    convolution applied twice.  Both uses run the native kernels (no layout bridge), the weight gradient is the sum
    over the uses, and the net remains eligible for a captured step."""
    body = """
layer[0->1] = fullc:enc
  nhidden = 64
layer[1->2] = tanh
layer[2->3] = share[enc]
layer[3->4] = relu
layer[4->5] = fullc:out
  nhidden = 7
layer[5->5] = softmax
"""
    errs, ref, b2 = compare_backends(_conf(body, 33, "1,1,64"), make_batch((33, 1, 1, 64), nclass=7), return_trainers=True)
    assert all(c.native for c in b2.net.connections), [c.native for c in b2.net.connections]
    assert b2.net.connections[2].layer is b2.net.connections[0].layer
    body = """
layer[0->1] = conv:stem
  kernel_size = 3
  pad = 1
  nchannel = 16
layer[1->2] = relu
layer[2->3] = conv:twice
  kernel_size = 3
  pad = 1
  nchannel = 16
layer[3->4] = relu
layer[4->5] = share[twice]
layer[5->6] = relu
layer[6->7] = flatten
layer[7->8] = fullc:out
  nhidden = 5
layer[8->8] = softmax
"""
    # (five bf16 layers deep on nine 10 x 10 images: rounding noise is several per cent of these small gradients; a
    # missing accumulation over the two uses would be an error of ~50 %)
    errs, ref, b2 = compare_backends(_conf(body, 9, "3,10,10"), make_batch((9, 3, 10, 10), nclass=5), tol_grad=0.12,
                                     return_trainers=True)
    assert errs["grad[2:wmat]"] == errs["grad[4:wmat]"]
    assert all(c.native for c in b2.net.connections), [c.native for c in b2.net.connections]
    # two updates through the public path: graphs stay possible (nothing forces the eager route)
    b2._graph_unsafe = None
    assert b2._graph_applicable(make_batch((9, 3, 10, 10), nclass=5))


def test_inner_strided_convolution_backpropagates_by_stride_phases():
    """a strided convolution that is not the first layer: its backward-data runs as stride-phase implicit GEMMs
    (ConvolutionLayer.dgrad_phases), no patch buffer is allocated, gradients match the reference backend"""
    body = """
layer[+1] = conv:stem
  kernel_size = 3
  pad = 1
  nchannel = 16
layer[+1] = relu
layer[+1] = conv:down
  kernel_size = 3
  pad = 1
  stride = 2
  nchannel = 32
layer[+1] = relu
layer[+1] = conv:down2
  kernel_size = 2
  stride = 2
  nchannel = 32
layer[+1] = flatten
layer[+1] = fullc:out
  nhidden = 6
layer[+0] = softmax
"""
    errs, ref, b2 = compare_backends(_conf(body, 10, "3,17,17"), make_batch((10, 3, 17, 17), nclass=6), tol_grad=0.08,
                                     return_trainers=True)
    down = [c.layer for c in b2.net.connections if c.layer.type_name == "conv"][1:]
    assert all(l.dgrad_phases and l.col is None for l in down)


@pytest.mark.parametrize("graph", [0, 1])
@pytest.mark.parametrize("updater", ["sgd", "nag", "adam"])
def test_updater_fused_into_weight_gradient_matches_separate_update(updater, graph, monkeypatch):
    """fully connected weight matrices on one GPU are updated inside their weight-gradient GEMM
    (cxn_gemm_wgrad_update) - same weights as GEMM + cxn_update after 4 steps of a scheduled learning rate, eagerly
    and through the captured step, also with gradient accumulation (where the fused path must stand down)"""
    from util import copy_weights, make_trainer
    body = """
layer[+1] = fullc:fc1
  nhidden = 96
layer[+1] = relu
layer[+0] = dropout
  threshold = 0.3
layer[+1] = fullc:fc2
  nhidden = 72
layer[+1] = tanh
layer[+1] = fullc:fc3
  nhidden = 10
layer[+0] = softmax
"""
    conf = (_conf(body, 32, "1,1,64") + "updater = %s\nwmat:wd = 0.001\nbias:lr = 0.2\nlr:schedule = expdecay\n"
            "lr:gamma = 0.5\nlr:step = 2\nclip_gradient = 0.5\ncuda_graph = %d\nseed = 11\n" % (updater, graph))
    for period in (1, 2):
        monkeypatch.setenv("CXN_FUSE_UPDATE", "0")
        plain = make_trainer(conf, dev="gpu:0", backend="b200", update_period=period)
        monkeypatch.setenv("CXN_FUSE_UPDATE", "1")
        fused = make_trainer(conf, dev="gpu:0", backend="b200", update_period=period)
        assert not plain.net.fuse_update and fused.net.fuse_update
        assert sum(s.fused for s in fused.net.arena.segments) == 3
        copy_weights(plain, fused)
        for s in range(4 * period):
            batch = make_batch((32, 1, 1, 64), seed=s, pin=True)
            plain.update(batch)
            fused.update(batch)
        assert fused.net.fuse_now == (period == 1)
        for name in ("fc1", "fc2", "fc3"):
            for tag in ("wmat", "bias"):
                wa, _ = plain.get_weight(name, tag)
                wb, _ = fused.get_weight(name, tag)
                assert torch.allclose(torch.from_numpy(wa), torch.from_numpy(wb), atol=1e-6, rtol=1e-5), \
                    (updater, graph, period, name, tag)
        # the bf16 copies the next forward pass reads follow the masters
        for s in fused.net.arena.segments:
            if s.shadow_off >= 0:
                sh = fused.net.arena.shadow_view(s).view(s.rows, s.shadow_ld)[:, :s.cols]
                assert torch.equal(sh, fused.net.arena.w_view(s).view(s.rows, s.cols).to(torch.bfloat16))


@pytest.mark.parametrize("cpad", ["1", "0"])
def test_conv_channel_padded_groups(cpad, monkeypatch, tmp_path):
    """48 channels per group (AlexNet conv2's shape class): the filter's internal layout pads every tap to 64 channels
    and each group runs over a channel window through im2col-mode TMA (CXN_CONV_CPAD, default on) - forward, weight
    gradient (the padding columns stay zero) and backward-data against the reference backend; the weights a user
    sees, and a checkpoint, keep the unpadded layout"""
    import io
    from util import copy_weights, make_trainer
    monkeypatch.setenv("CXN_CONV_CPAD", cpad)
    body = """
layer[+1] = conv:c1
  kernel_size = 3
  pad = 1
  nchannel = 96
layer[+1] = relu
layer[+1] = conv:c2
  kernel_size = 5
  pad = 2
  ngroup = 2
  nchannel = 64
layer[+1] = relu
layer[+1] = conv:c3
  kernel_size = 3
  pad = 1
  nchannel = 48
layer[+1] = relu
layer[+1] = conv:c4
  kernel_size = 3
  nchannel = 16
layer[+1] = flatten
layer[+1] = fullc:fc1
  nhidden = 10
layer[+0] = softmax
"""
    conf = _conf(body, 8, "8,11,10") + "eta = 0.01\nmomentum = 0.5\n"      # small steps: three updates stay comparable
    _, ref, b2 = compare_backends(conf, make_batch((8, 8, 11, 10)), tol_grad=0.08, return_trainers=True)
    convs = {c.layer.layer_index: c.layer for c in b2.net.connections if type(c.layer).__name__ == "ConvolutionLayer"}
    pads = sorted(l.cpad for l in convs.values())
    assert pads == ([0, 0, 64, 64] if cpad == "1" else [0, 0, 0, 0])       # c2 (2 x 48) and c4 (1 x 48)
    # three training steps: the structural zeros stay zero; the visible weights follow the reference backend (bf16
    # drift of a small deep net) and, closely, the same net in the other filter layout
    monkeypatch.setenv("CXN_CONV_CPAD", "0" if cpad == "1" else "1")
    twin = make_trainer(conf, dev="gpu:0", backend="b200")
    monkeypatch.setenv("CXN_CONV_CPAD", cpad)
    copy_weights(ref, b2)
    copy_weights(ref, twin)
    for s in range(3):
        batch = make_batch((8, 8, 11, 10), seed=10 + s)
        ref.update(batch)
        b2.update(batch)
        twin.update(batch)
    for name in ("c2", "c4"):
        wr, _ = ref.get_weight(name, "wmat")
        wb, shape = b2.get_weight(name, "wmat")
        wt, _ = twin.get_weight(name, "wmat")
        assert wb.shape == wr.shape
        assert rel_err_np(wr, wb) < 2e-2, name
        assert rel_err_np(wt, wb) < 2e-2, name
    if cpad == "1":
        for l in convs.values():
            if l.cpad:
                p = l.param
                cin = p.num_input_channel // p.num_group
                w5 = l.bind["wmat"].w.view(-1, p.kernel_height, p.kernel_width, l.cpad)
                assert float(w5[..., cin:].abs().max()) == 0.0
                g5 = l.bind["wmat"].g.view(-1, p.kernel_height, p.kernel_width, l.cpad)
                assert float(g5[..., cin:].abs().max()) == 0.0
    # checkpoint round trip into the other layout
    buf = io.BytesIO()
    b2.save_model(buf)
    monkeypatch.setenv("CXN_CONV_CPAD", "0" if cpad == "1" else "1")
    other = make_trainer(conf, dev="gpu:0", backend="b200")
    buf.seek(0)
    other.load_model(buf)
    for name in ("c2", "c4"):
        wa, _ = b2.get_weight(name, "wmat")
        wo, _ = other.get_weight(name, "wmat")
        assert (wa == wo).all()


def rel_err_np(a, b):
    import numpy as np
    return float(np.abs(a - b).sum() / max(np.abs(a).sum(), 1e-12))


@pytest.mark.parametrize("k,shape", [(11, (3, 67, 67)), (8, (3, 64, 60)), (5, (1, 45, 43)), (11, (3, 227, 227))])
def test_conv_first_layer_space_to_depth(k, shape, monkeypatch):
    """stride-4 first layer over the space-to-depth input node ([N][H/4][W/4][64]): a stride-1 convolution with a
    ceil(k/4)^2 filter through the ordinary im2col-TMA kernels - forward, weight gradient (taps outside the k x k
    window stay zero), input gradient and the layout bridge of node 0, against the reference backend; 8-bit and bf16
    batches convert to the same node contents as fp32 batches"""
    body = """
layer[+1] = conv:cv1
  kernel_size = %d
  stride = 4
  nchannel = 24
layer[+1] = relu
layer[+1] = flatten
layer[+1] = fullc:fc1
  nhidden = 10
layer[+0] = softmax
""" % k
    nb = 5 if shape[1] > 100 else 9
    shp = "%d,%d,%d" % shape
    errs, ref, b2 = compare_backends(_conf(body, nb, shp), make_batch((nb,) + shape), return_trainers=True, tol_grad=0.1)
    layer = b2.net.connections[0].layer
    assert layer.s2d and not layer.span and not layer.pair
    node0 = b2.net.nodes[0]
    assert node0.s2d == 4 and node0.act.ld == 64
    # the structural zeros collected no gradient, the real taps did
    k3 = (k + 3) // 4
    g = layer.bind["wmat"].g.reshape(24, k3, k3, 4, 4, 4)          # (oc, ky, kx, dy, dx, c)
    m = layer.s2d_mask.reshape(k3, k3, 4, 4, 4)
    assert float((g * (1 - m)).abs().max()) == 0.0
    assert float((g * m).abs().max()) > 0.0
    assert int(m.sum()) == k * k * shape[0]
    # node 0 through the bridge, and the 8-bit / bf16 conversions
    import torch as T
    from cxxnet_b200 import ops
    x = make_batch((nb,) + shape).data
    got = b2.net._node_load(node0)
    assert T.equal(got.cpu(), x.to(T.bfloat16).float())
    keep = node0.act.t.clone()
    L = ops.lib()
    xb = x.to(T.bfloat16).cuda()
    node0.act.t.zero_()
    ops.check(L.cxn_nchw_to_s2d4(xb.data_ptr(), 1, node0.act.ptr, nb, shape[0], shape[1], shape[2], 0.0, 0.0, 0.0, 0.0, 1.0,
                                 ops.stream()), "s2d bf16")
    assert T.equal(node0.act.t, keep)
    u8 = T.randint(0, 256, (nb,) + shape, dtype=T.uint8)
    mean = [100.0, 110.0, 120.0, 0.0]
    ops.check(L.cxn_nchw_to_s2d4(u8.cuda().data_ptr(), 2, node0.act.ptr, nb, shape[0], shape[1], shape[2], mean[0], mean[1],
                                 mean[2], mean[3], 0.01, ops.stream()), "s2d u8")
    want = ((u8.float() - T.tensor(mean[:shape[0]]).view(1, -1, 1, 1)) * 0.01).to(T.bfloat16).float()
    assert T.equal(b2.net._node_load(node0).cpu(), want)
    # weights a user sees keep the reference layout
    w, shape_w = b2.get_weight("cv1", "wmat")
    wr, _ = ref.get_weight("cv1", "wmat")
    assert w.shape == wr.shape
