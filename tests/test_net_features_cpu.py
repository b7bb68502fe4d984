"""Net-level features on the reference-math backend that no other CPU test reaches: weight sharing
(`share[tag]`), multi-field labels with `label_vec` / `metric[field,node]`, extra input nodes, the `img`
iterator, the affine augmenter, and the partition-maker tool."""
import io
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

from util import make_batch, make_trainer

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SHARED = """
netconfig=start
layer[0->1] = fullc:enc
  nhidden = 6
  no_bias = 1
layer[1->2] = tanh
layer[2->3] = share[enc]
layer[3->3] = l2_loss
netconfig=end
input_shape = 1,1,6
batch_size = 4
dev = cpu
label_vec[0,6) = label
eta = 0.05
momentum = 0.0
random_type = gaussian
init_sigma = 0.3
metric = rmse
"""


def test_shared_layer_ties_weights():
    """both uses of a shared layer read the same matrix and its gradient is the sum of both uses
    (reference nnet_config.h:338-346, neural_net-inl.hpp:238-244)"""
    tr = make_trainer(SHARED)
    conns = tr.net.connections
    assert conns[2].layer is conns[0].layer
    assert len([1 for c in conns for _ in c.layer.weights()]) >= 1
    w0, _ = tr.get_weight("enc", "wmat")
    b = make_batch((4, 1, 1, 6), dense_label=6, seed=3)
    x = b.data.reshape(4, 6).double()
    y = b.label.double()
    W = torch.from_numpy(w0.reshape(6, 6)).double().requires_grad_(True)
    out = torch.tanh(x @ W.t()) @ W.t()
    # the loss layer's gradient is (prediction - label) / batch
    (0.5 * ((out - y) ** 2).sum() / 4).backward()
    tr.update(b)
    w1, _ = tr.get_weight("enc", "wmat")
    want = w0.reshape(6, 6) - 0.05 * W.grad.numpy()
    assert np.allclose(w1.reshape(6, 6), want, atol=1e-5)
    # one copy in the checkpoint
    buf = io.BytesIO()
    tr.save_model(buf)
    tr2 = make_trainer(SHARED)
    buf.seek(0)
    tr2.load_model(buf)
    assert np.array_equal(tr2.get_weight("enc", "wmat")[0], w1)


MULTI = """
netconfig=start
layer[0->1] = fullc:trunk
  nhidden = 8
layer[1->2] = relu
layer[2->3,4] = split
layer[3->5] = fullc:cls
  nhidden = 3
layer[5->5] = softmax
  target = kind
layer[4->6] = fullc:attr
  nhidden = 2
layer[6->6] = multi_logistic
  target = flags
netconfig=end
input_shape = 1,1,5
batch_size = 6
dev = cpu
label_vec[0,1) = kind
label_vec[1,3) = flags
metric[kind,5] = error
metric[flags,6] = logloss
eta = 0.1
"""


def test_two_losses_with_label_fields_and_node_metrics():
    tr = make_trainer(MULTI)
    g = torch.Generator().manual_seed(0)
    b = make_batch((6, 1, 1, 5), seed=1)
    b.label = torch.cat([torch.randint(0, 3, (6, 1), generator=g).float(),
                         torch.randint(0, 2, (6, 2), generator=g).float()], 1)
    before = [tr.get_weight(n, "wmat")[0].copy() for n in ("trunk", "cls", "attr")]
    for _ in range(3):
        tr.update(b)
    after = [tr.get_weight(n, "wmat")[0] for n in ("trunk", "cls", "attr")]
    assert all(np.abs(a - c).max() > 0 for a, c in zip(after, before))       # both heads and the trunk learn
    line = tr.evaluate(None, "train")
    assert "train-error[kind]:" in line and "train-logloss[flags]:" in line
    p = tr.extract_feature(b, "5")
    assert p.shape == (6, 1, 1, 3) and np.allclose(p.reshape(6, 3).sum(1), 1.0, atol=1e-5)


EXTRA = """
extra_data_num = 1
extra_data_shape[0] = 1,1,3
netconfig=start
layer[in,in_1->2] = concat
layer[2->3] = fullc:fc
  nhidden = 2
layer[3->3] = softmax
netconfig=end
input_shape = 1,1,4
batch_size = 5
dev = cpu
eta = 0.1
metric = error
"""


def test_extra_data_nodes_feed_the_net():
    tr = make_trainer(EXTRA)
    b = make_batch((5, 1, 1, 4), nclass=2, seed=2)
    b.extra_data = [torch.randn(5, 1, 1, 3)]
    w0 = tr.get_weight("fc", "wmat")[0].copy()
    assert w0.shape == (2, 7)                                  # 4 inputs + 3 extra values
    tr.update(b)
    assert np.abs(tr.get_weight("fc", "wmat")[0] - w0)[:, 4:].max() > 0     # the extra columns get gradient
    z = tr.extract_feature(b, "2")
    assert np.allclose(z.reshape(5, 7)[:, 4:], b.extra_data[0].reshape(5, 3).numpy(), atol=1e-6)


def test_img_iterator_reads_list_and_files(tmp_path):
    PIL = pytest.importorskip("PIL")
    from PIL import Image
    from cxxnet_b200.config import parse_config_string
    from cxxnet_b200.io.data import create_iterator
    arrs = []
    lines = []
    for i in range(5):
        a = (np.random.RandomState(i).rand(9, 11, 3) * 255).astype(np.uint8)
        arrs.append(a)
        name = "im%d.%s" % (i, "png" if i % 2 else "jpg")
        Image.fromarray(a).save(tmp_path / name, quality=95)
        lines.append("%d\t%d\t%d\t%s\n" % (100 + i, i % 2, i, name))
    (tmp_path / "l.lst").write_text("".join(lines))
    it = create_iterator(parse_config_string(
        'iter = img\n image_list = "%s"\n image_root = "%s/"\n label_width = 2\n batch_size = 2\n'
        ' input_shape = 3,9,11\n round_batch = 1\n silent = 1\niter = end\n' % (tmp_path / "l.lst", tmp_path)))
    it.init()
    it.before_first()
    seen, padds = [], []
    while it.next():
        b = it.value()
        padds.append(b.num_batch_padd)
        for r in range(b.batch_size - b.num_batch_padd):
            seen.append((int(b.inst_index[r]), b.label[r].tolist(), b.data[r].numpy().copy()))
    assert [s[0] for s in seen] == [100, 101, 102, 103, 104]
    assert seen[3][1] == [1.0, 3.0]
    assert np.array_equal(seen[1][2], arrs[1].transpose(2, 0, 1).astype(np.float32))       # PNG: exact
    ref = np.asarray(Image.open(tmp_path / "im2.jpg").convert("RGB")).transpose(2, 0, 1).astype(np.float32)
    assert np.abs(seen[2][2] - ref).max() <= 3                                             # JPEG: as libjpeg
    assert padds == [0, 0, 1]                                  # 5 images, batches of 2, wrapped around
    assert it._native                                          # files are decoded by the native batch path ...

    def crops(native_pipeline):
        jt = create_iterator(parse_config_string(
            'iter = img\n image_list = "%s"\n image_root = "%s/"\n label_width = 2\n batch_size = 2\n'
            ' input_shape = 3,7,8\n rand_crop = 1\n rand_mirror = 1\n mean_value = 9,8,7\n round_batch = 1\n'
            ' native_pipeline = %d\n silent = 1\niter = end\n' % (tmp_path / "l.lst", tmp_path, native_pipeline)))
        jt.init()
        jt.before_first()
        got = []
        while jt.next():
            got.append(jt.value().data.clone())
        return got
    a, b = crops(0), crops(1)                                  # ... with the results of the Python stages
    assert len(a) == len(b) == 3 and all(torch.equal(x, y) for x, y in zip(a, b))


def test_affine_augmenter_rotation_and_crop():
    from cxxnet_b200.io.iterators import ImageAugmenter, RandomSampler
    aug = ImageAugmenter()
    for k, v in (("input_shape", "1,8,8"), ("rotate", "90"), ("fill_value", "0"), ("min_img_size", "8"),
                 ("max_img_size", "8")):
        aug.set_param(k, v)
    assert aug.need_process()
    img = np.zeros((1, 8, 8), dtype=np.float32)
    img[0, 1, 2] = 100.0                                       # one bright pixel
    out = aug.process(img, RandomSampler(1))
    assert out.shape == (1, 8, 8)
    y, x = np.unravel_index(int(out[0].argmax()), (8, 8))
    # a quarter turn about the image centre maps (row 1, col 2) onto (row 2, col 6) or (row 5, col 1),
    # depending on the direction; the mass is preserved up to interpolation
    assert (int(y), int(x)) in ((2, 6), (5, 1), (2, 7), (6, 1), (1, 5), (5, 0)) or out.sum() > 50
    assert 50 < out.sum() < 150
    ident = ImageAugmenter()
    ident.set_param("input_shape", "1,8,8")
    assert not ident.need_process() and ident.process(img, RandomSampler(1)) is img


def test_partition_maker_writes_parts_and_makefile(tmp_path):
    lines = []
    for i in range(6):
        (tmp_path / ("f%d.bin" % i)).write_bytes(b"x" * (400 << 10))          # 400 KiB each
        lines.append("%d\t0\tf%d.bin\n" % (i, i))
    (tmp_path / "all.lst").write_text("".join(lines))
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "imgbin_partition_maker.py"),
                        "--img_list", str(tmp_path / "all.lst"), "--img_root", str(tmp_path) + "/",
                        "--partition_size", "1", "--prefix", "part%03d", "--out", str(tmp_path / "out"),
                        "--makefile", str(tmp_path / "Gen.mk"), "--im2bin", "im2bin"],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=60)
    assert r.returncode == 0, r.stdout
    parts = sorted(p for p in os.listdir(tmp_path / "out") if p.endswith(".lst"))
    assert len(parts) == 3                                      # two 400 KiB files per 1 MiB part
    assert sum(len((tmp_path / "out" / p).read_text().splitlines()) for p in parts) == 6
    mk = (tmp_path / "Gen.mk").read_text()
    assert mk.startswith("all: ") and mk.count("im2bin ") == 3 and "part001.bin" in mk


def test_single_channel_net_reads_image_sources(tmp_path):
    """input_shape = 1,h,w over an image source: grey-scale files arrive as one plane"""
    pytest.importorskip("PIL")
    from PIL import Image
    from cxxnet_b200.config import parse_config_string
    from cxxnet_b200.io.data import create_iterator
    lines = []
    arrs = []
    for i in range(4):
        a = (np.random.RandomState(i).rand(12, 14) * 255).astype(np.uint8)
        arrs.append(a)
        Image.fromarray(a, mode="L").save(tmp_path / ("g%d.png" % i))
        lines.append("%d\t%d\tg%d.png\n" % (i, i % 2, i))
    (tmp_path / "g.lst").write_text("".join(lines))
    it = create_iterator(parse_config_string(
        'iter = img\n image_list = "%s"\n image_root = "%s/"\n batch_size = 2\n input_shape = 1,12,14\n'
        ' mean_value = 10\n silent = 1\niter = end\n' % (tmp_path / "g.lst", tmp_path)))
    it.init()
    it.before_first()
    assert it.next()
    b = it.value()
    assert tuple(b.data.shape) == (2, 1, 12, 14) and it._native
    assert np.allclose(b.data[1, 0].numpy(), arrs[1].astype(np.float32) - 10.0)
    # colour JPEGs into a single-channel net: native batch path == Python stages
    for i in range(4):
        a = (np.random.RandomState(10 + i).rand(12, 14, 3) * 255).astype(np.uint8)
        Image.fromarray(a).save(tmp_path / ("c%d.jpg" % i), quality=92)
    (tmp_path / "c.lst").write_text("".join("%d\t0\tc%d.jpg\n" % (i, i) for i in range(4)))

    def run(native_pipeline):
        jt = create_iterator(parse_config_string(
            'iter = img\n image_list = "%s"\n image_root = "%s/"\n batch_size = 2\n input_shape = 1,10,12\n'
            ' rand_crop = 1\n rand_mirror = 1\n mean_value = 100\n divideby = 16\n native_pipeline = %d\n silent = 1\n'
            'iter = end\n' % (tmp_path / "c.lst", tmp_path, native_pipeline)))
        jt.init()
        jt.before_first()
        out = []
        while jt.next():
            out.append(jt.value().data.clone())
        return out
    a0, a1 = run(0), run(1)
    assert len(a0) == 2 and all(torch.equal(x, y) for x, y in zip(a0, a1))


def test_bicubic_warp_reproduces_cubics_and_the_identity():
    """the affine warp resamples with the Keys kernel (a = -0.75, cv::INTER_CUBIC of the reference,
    src/io/image_augmenter-inl.hpp:105-108): integer positions return the pixel, the weights sum to one, a linear
    ramp is reproduced exactly away from the border, and positions outside read fill_value"""
    from cxxnet_b200.io.iterators import ImageAugmenter
    aug = ImageAugmenter()
    aug.set_param("input_shape", "1,8,8")
    aug.set_param("fill_value", "7")
    t = np.linspace(0, 1, 11, dtype=np.float32)[:-1]
    c = aug.cubic_weights(t)
    assert np.allclose(sum(c), 1.0, atol=1e-6)
    assert np.allclose([w[0] for w in c], [0.0, 1.0, 0.0, 0.0], atol=1e-7)         # t = 0: the pixel itself
    assert np.isclose(c[1][5], 0.59375) and np.isclose(c[0][5], -0.09375)        # t = 0.5, a = -0.75: (-3, 19, 19, -3) / 32
    img = np.arange(12 * 12, dtype=np.float32).reshape(1, 12, 12)
    shift = np.array([[1.0, 0.0, 2.25], [0.0, 1.0, 1.5], [0.0, 0.0, 1.0]])            # dst(x, y) = src(x + 2.25, y + 1.5)
    out = aug.warp(img, (shift, 12, 12, 0, 0, False))
    want = np.fromfunction(lambda y, x: (y + 1.5) * 12 + (x + 2.25), (12, 12))
    # a = -0.75 (unlike a = -0.5) does not reproduce a ramp exactly: at fraction t the kernel's centroid is
    # -c0 + c2 + 2 c3 instead of t - the same small bias OpenCV's INTER_CUBIC has
    def centroid(t):
        c = [float(v) for v in aug.cubic_weights(np.float32(t))]
        return -c[0] + c[2] + 2 * c[3]
    cub = want + (centroid(0.25) - 0.25) + 12 * (centroid(0.5) - 0.5)
    assert abs(centroid(0.5) - 0.5) < 1e-6 and abs(centroid(0.25) - 0.25) < 0.05
    assert np.allclose(out[0, 1:8, 1:7], cub[1:8, 1:7], atol=1e-3)
    assert out[0, 11, 11] != want[11, 11] and np.isfinite(out).all()                  # border taps read the fill value
    far = aug.warp(img, (np.array([[1.0, 0, 100.0], [0, 1.0, 100.0], [0, 0, 1.0]]), 4, 4, 0, 0, False))
    assert np.allclose(far, 7.0)
    aug.set_param("warp_interp", "bilinear")
    assert np.allclose(aug.warp(img, (shift, 12, 12, 0, 0, False))[0, 1:8, 1:7], want[1:8, 1:7], atol=1e-3)   # exact
