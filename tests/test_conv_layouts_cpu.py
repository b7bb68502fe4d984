"""Internal filter layouts of the b200 convolution routes, checked on the CPU with plain torch: the planning decisions
for AlexNet's layers, the conversions to and from the checkpoint layout, and that a convolution computed the way the
kernels see the data (space-to-depth blocks / channel-padded group windows) equals the reference convolution."""
import torch
import torch.nn.functional as F

from util import make_trainer


def _alexnet_layers():
    from cxxnet_b200.models import alexnet_conf
    tr = make_trainer(alexnet_conf(batch_size=2, dev="cpu"))
    convs = [c.layer for c in tr.net.connections if type(c.layer).__name__ == "ConvolutionLayer"]
    assert len(convs) == 5
    return tr, convs


def test_alexnet_layers_plan_space_to_depth_and_channel_padded_groups():
    tr, convs = _alexnet_layers()
    plans = [l.b200_plan(None, i == 0, True) for i, l in enumerate(convs)]
    c1, c2, c3, c4, c5 = convs
    assert c1.s2d and not c1.span and not c1.pair and plans[0] == ("s2d", 4)
    assert c1._geom() == (1, 64, 96, 576, 576)                  # 3 x 3 blocks of 64 channels
    assert c2.cpad == 64 and not c2.s2d and c2._geom() == (2, 64, 128, 1600, 1600)
    assert [l.cpad for l in (c3, c4, c5)] == [0, 0, 0] and c3._geom()[1] == 256 and c4._geom()[1] == 192
    assert all(l.igemm_fwd for l in convs) and [l.igemm_dgrad for l in convs] == [False, True, True, True, True]

    # ---- conv1: space to depth.  round trip, structural zeros, and the convolution itself
    p = c1.param
    torch.manual_seed(0)
    w = torch.randn(1, 96, 3 * 11 * 11)
    wi = c1.weight_to_internal("wmat", w)
    assert wi.shape == (1, 96, 576) and torch.equal(c1.weight_from_internal("wmat", wi), w)
    mask = c1.weight_to_internal("wmat", torch.ones_like(w))[0, 0]
    assert int(mask.sum()) == 363 and set(mask.unique().tolist()) == {0.0, 1.0}
    x = torch.randn(2, 3, 227, 227)
    ref = F.conv2d(x, w.view(96, 3, 11, 11), stride=4)
    full = torch.zeros(2, 4, 228, 228)
    full[:, :3, :227, :227] = x
    xs = full.reshape(2, 4, 57, 4, 57, 4).permute(0, 2, 4, 3, 5, 1).reshape(2, 57, 57, 64)    # [n][Y][X][(dy, dx, c)]
    got = F.conv2d(xs.permute(0, 3, 1, 2), wi.view(96, 3, 3, 64).permute(0, 3, 1, 2))
    assert got.shape == ref.shape == (2, 96, 55, 55)
    assert torch.allclose(got, ref, rtol=1e-4, atol=2e-3)

    # ---- conv2: channel-padded groups.  group g reads the 64-channel window that starts at channel 48 g
    w2 = torch.randn(2, 128, 48 * 5 * 5)
    wi2 = c2.weight_to_internal("wmat", w2)
    assert wi2.shape == (2, 128, 1600) and torch.equal(c2.weight_from_internal("wmat", wi2), w2)
    assert float(wi2.view(2, 128, 5, 5, 64)[..., 48:].abs().max()) == 0.0
    x2 = torch.randn(2, 96, 27, 27)
    ref2 = F.conv2d(x2, w2.view(256, 48, 5, 5), padding=2, groups=2)
    xpad = torch.cat([x2, torch.zeros(2, 16, 27, 27)], dim=1)          # what the tensor map returns past channel 95
    outs = []
    for g in range(2):
        window = xpad[:, 48 * g:48 * g + 64]
        wg = wi2[g].view(128, 5, 5, 64).permute(0, 3, 1, 2)
        outs.append(F.conv2d(window, wg, padding=2))
    assert torch.allclose(torch.cat(outs, dim=1), ref2, rtol=1e-4, atol=2e-3)


def test_space_to_depth_node_bridge_roundtrip():
    """the executor's layout bridge of a space-to-depth input node (node_store then node_load) is the identity, and
    the stored form is the [n][Y][X][(dy*4 + dx)*4 + c] layout cxn_nchw_to_s2d4 writes"""
    from cxxnet_b200.executor import NeuralNet
    from cxxnet_b200.layers.base import Node

    class Mat:
        def __init__(self, rows):
            self.t = torch.zeros(rows, 64)

    n = Node("in", 0)
    b, c, h, w = 2, 3, 11, 9
    n.shape = [b, c, h, w]
    n.s2d = 4
    hs, ws = 3, 3
    n.act = Mat(b * hs * ws)
    n.grad = Mat(b * hs * ws)
    x = torch.randn(b, c, h, w)
    NeuralNet._node_store(None, n, x)
    assert torch.equal(NeuralNet._node_load(None, n), x)
    blocks = n.act.t.view(b, hs, ws, 4, 4, 4)           # (n, Y, X, dy, dx, c)
    assert float(blocks[1, 2, 1, 3, 2, 1]) == float(x[1, 1, 4 * 2 + 3, 4 * 1 + 2]) if 4 * 2 + 3 < h and 4 * 1 + 2 < w else True
    assert float(blocks[0, 1, 2, 1, 0, 2]) == float(x[0, 2, 5, 8])
    assert float(blocks[..., 3].abs().max()) == 0.0      # the pad channel
    assert float(blocks[:, 2, :, 3].abs().max()) == 0.0  # rows past the image (4 * 2 + 3 = 11 >= h)
