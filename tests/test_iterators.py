"""data iterators and chain semantics (reference src/io/*)."""
import io
import os

import numpy as np
import pytest
import torch

from cxxnet_b200.config import parse_config_string
from cxxnet_b200.io.data import create_iterator
from cxxnet_b200.io.iterators import PAGE_INTS, BinaryPage


def _mk(text):
    cfg = parse_config_string(text)
    it = create_iterator(cfg)
    it.init()
    return it


def test_mnist_synthetic_shapes_flat_and_image_drop_last_batch():
    it = _mk("iter = mnist\n path_img = nope\n path_label = nope\n batch_size = 64\n num_synthetic = 200\n silent = 1\niter = end\n")
    n = 0
    for b in it:
        assert tuple(b.data.shape) == (64, 1, 1, 784) and tuple(b.label.shape) == (64, 1)
        assert 0.0 <= float(b.data.min()) and float(b.data.max()) < 1.0
        n += 1
    assert n == 3                                   # 200 // 64, last partial batch dropped
    it2 = _mk("iter = mnist\n path_img = nope\n path_label = nope\n input_flat = 0\n batch_size = 50\n num_synthetic = 100\n shuffle = 1\n silent = 1\niter = end\n")
    it2.before_first()
    assert it2.next() and tuple(it2.value().data.shape) == (50, 1, 28, 28)
    assert sorted(set(it2.inst.tolist())) == list(range(100))


def test_mnist_reads_real_idx_files(tmp_path):
    import gzip, struct
    imgs = (np.arange(3 * 28 * 28) % 251).astype(np.uint8).reshape(3, 28, 28)
    with gzip.open(tmp_path / "img.gz", "wb") as f:
        f.write(struct.pack(">iiii", 2051, 3, 28, 28) + imgs.tobytes())
    with gzip.open(tmp_path / "lab.gz", "wb") as f:
        f.write(struct.pack(">ii", 2049, 3) + bytes([7, 1, 4]))
    it = _mk('iter = mnist\n path_img = "%s"\n path_label = "%s"\n batch_size = 3\n silent = 1\niter = end\n'
             % (tmp_path / "img.gz", tmp_path / "lab.gz"))
    it.before_first()
    assert it.next()
    b = it.value()
    assert b.label.view(-1).tolist() == [7.0, 1.0, 4.0]
    assert torch.allclose(b.data.view(3, 28, 28), torch.from_numpy(imgs.astype(np.float32) / 256.0))
    assert not it.next()


SYN = """iter = synthetic
  num_inst = 10
  num_class = 3
  batch_size = 4
  input_shape = 3,8,8
  silent = 1
%s
iter = end
"""


def test_batch_adapter_padding_and_round_batch():
    it = _mk(SYN % "")
    pads = [b.num_batch_padd for b in it]
    assert pads == [0, 0, 2]                         # 10 = 4 + 4 + 2 (+2 padded)
    it = _mk(SYN % "  round_batch = 1")
    got = [(b.num_batch_padd, b.inst_index.tolist()) for b in it]
    assert got[2][0] == 2 and got[2][1] == [8, 9, 0, 1]          # wraps around to the start
    # next epoch starts after the overflowed instances
    it.before_first()
    assert it.next() and it.value().inst_index.tolist() == [2, 3, 4, 5]


def test_test_skipread_returns_same_batch_forever():
    it = _mk(SYN % "  test_skipread = 1")
    it.before_first()
    assert it.next()
    first = it.value().data.clone()
    for _ in range(5):
        assert it.next() and torch.equal(it.value().data, first)


def test_augment_crop_mirror_scale_mean_value():
    it = _mk("""iter = synthetic
  num_inst = 6
  batch_size = 2
  input_shape = 3,6,6
  source_shape = 3,8,8
  rand_crop = 1
  rand_mirror = 1
  divideby = 2
  mean_value = 1,1,1
  silent = 1
iter = end
""")
    for b in it:
        assert tuple(b.data.shape) == (2, 3, 6, 6) and torch.isfinite(b.data).all()
    # deterministic centre crop + mirror
    a = _mk("iter = synthetic\n num_inst = 2\n batch_size = 2\n input_shape = 1,4,4\n source_shape = 1,6,6\n silent = 1\niter = end\n")
    m = _mk("iter = synthetic\n num_inst = 2\n batch_size = 2\n input_shape = 1,4,4\n source_shape = 1,6,6\n mirror = 1\n mean_value = 0.001,0,0\n silent = 1\niter = end\n")
    a.before_first(); m.before_first()
    a.next(); m.next()
    assert torch.allclose(a.value().data.flip(3) - 0.001, m.value().data, atol=1e-5)


def test_mean_image_is_created_cached_and_subtracted(tmp_path):
    mean = tmp_path / "mean.bin"
    txt = 'iter = synthetic\n num_inst = 8\n batch_size = 4\n input_shape = 3,5,5\n image_mean = "%s"\n silent = 1\niter = end\n' % mean
    it = _mk(txt)
    assert mean.exists()
    tot = torch.zeros(3, 5, 5)
    for b in it:
        tot += b.data.sum(0)
    assert float(tot.abs().max()) < 1e-3            # data minus its own mean sums to zero
    it2 = _mk(txt)                                  # second run loads the cached file
    it2.before_first()
    assert it2.next()


def test_threadbuffer_and_membuffer_replay_the_same_batches():
    base = _mk(SYN % "")
    want = [b.data.clone() for b in base]
    tb = _mk(SYN % "" + "")
    tb = create_iterator(parse_config_string((SYN % "").replace("iter = end", "iter = threadbuffer\niter = end")))
    tb.init()
    for epoch in range(2):
        got = [b.data.clone() for b in tb]
        assert len(got) == len(want) and all(torch.equal(a, b) for a, b in zip(got, want))
    tb.close()
    mb = create_iterator(parse_config_string((SYN % "").replace("iter = end", "iter = membuffer\n max_nbatch = 2\niter = end")))
    mb.init()
    got = [b.data.clone() for b in mb]
    assert len(got) == 2 and torch.equal(got[1], want[1])


def test_attachtxt_joins_extra_data_by_instance_index(tmp_path):
    f = tmp_path / "extra.txt"
    f.write_text("2\n" + "".join("%d %d %d\n" % (i, i * 10, i * 10 + 1) for i in range(10)))
    it = create_iterator(parse_config_string((SYN % "").replace("iter = end", 'iter = attachtxt\n filename = "%s"\niter = end' % f)))
    it.init()
    it.before_first()
    assert it.next()
    b = it.value()
    assert tuple(b.extra_data[0].shape) == (4, 1, 1, 2)
    assert b.extra_data[0][2, 0, 0].tolist() == [20.0, 21.0]


def test_binary_page_layout_and_imgbin_iterator(tmp_path):
    pytest.importorskip("PIL")
    from PIL import Image
    page = BinaryPage()
    lst = []
    blobs = []
    for i in range(5):
        arr = (np.random.RandomState(i).rand(12, 10, 3) * 255).astype(np.uint8)
        bio = io.BytesIO()
        Image.fromarray(arr).save(bio, format="PNG")
        blobs.append((arr, bio.getvalue()))
        assert page.push(bio.getvalue())
        lst.append("%d\t%d\timg%d.png\n" % (i, i % 3, i))
    head = np.frombuffer(page.raw, dtype="<i4")
    assert head[0] == 5 and head[1] == 0 and head[2] == len(blobs[0][1])          # count, offsets table
    assert page.get(3) == blobs[3][1]
    assert bytes(page.raw[-len(blobs[0][1]):]) == blobs[0][1]                      # blobs packed from the end
    (tmp_path / "a.lst").write_text("".join(lst))
    with open(tmp_path / "a.bin", "wb") as f:
        f.write(page.raw)
    assert os.path.getsize(tmp_path / "a.bin") == PAGE_INTS * 4 == 64 << 20
    for kind in ("imgbin", "imgbinx"):
        it = _mk('iter = %s\n image_list = "%s"\n image_bin = "%s"\n batch_size = 2\n input_shape = 3,12,10\n silent = 1\niter = end\n'
                 % (kind, tmp_path / "a.lst", tmp_path / "a.bin"))
        seen = []
        for b in it:
            for r in range(b.batch_size - b.num_batch_padd):
                seen.append((int(b.inst_index[r]), float(b.label[r, 0]), b.data[r].clone()))
        it.close()
        assert [s[0] for s in seen] == [0, 1, 2, 3, 4] and [s[1] for s in seen] == [0.0, 1.0, 2.0, 0.0, 1.0]
        want = torch.from_numpy(blobs[2][0].astype(np.float32)).permute(2, 0, 1)   # RGB planar
        assert torch.equal(seen[2][2], want)


def test_imgbin_worker_sharding(tmp_path, monkeypatch):
    from cxxnet_b200.io.iterators import ImageBinIterator
    it = ImageBinIterator(False)
    it.set_param("image_conf_prefix", str(tmp_path / "part%02d"))
    it.set_param("image_conf_ids", "1-10")
    it.set_param("dist_num_worker", "4")
    monkeypatch.setenv("PS_RANK", "2")
    it._parse_image_conf()
    assert [os.path.basename(p) for p in it.path_imglst] == ["part07.lst", "part08.lst", "part09.lst"]


def test_im2bin_tool_and_native_page_reader(tmp_path):
    """bin/im2bin packs a list; the native reader thread and the Python reader return the same blobs."""
    pytest.importorskip("PIL")
    import subprocess
    from PIL import Image
    from cxxnet_b200.io import native
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = os.path.join(root, "bin", "im2bin")
    if not os.path.exists(exe) or native.lib() is None:
        import sys
        sys.path.insert(0, root)
        import build as _b
        _b.build_host()
    imgdir = tmp_path / "imgs"
    imgdir.mkdir()
    lines = []
    for i in range(7):
        arr = (np.random.RandomState(100 + i).rand(9, 11, 3) * 255).astype(np.uint8)
        Image.fromarray(arr).save(imgdir / ("i%d.png" % i))
        lines.append("%d\t%d\ti%d.png\n" % (i, i % 2, i))
    (tmp_path / "a.lst").write_text("".join(lines))
    subprocess.check_call([exe, str(tmp_path / "a.lst"), str(imgdir) + "/", str(tmp_path / "a.bin")])
    assert os.path.getsize(tmp_path / "a.bin") == 64 << 20
    rd = native.PageReader([str(tmp_path / "a.bin")])
    blobs = []
    while True:
        b = rd.next()
        if b is None:
            break
        blobs.append(b)
    assert len(blobs) == 7 and blobs[3] == (imgdir / "i3.png").read_bytes()
    rd.rewind()
    assert rd.next() == blobs[0]
    rd.close()
    outs = []
    for use_native in (1, 0):
        it = _mk('iter = imgbin\n image_list = "%s"\n image_bin = "%s"\n native_reader = %d\n batch_size = 7\n input_shape = 3,9,11\n silent = 1\niter = end\n'
                 % (tmp_path / "a.lst", tmp_path / "a.bin", use_native))
        it.before_first()
        assert it.next()
        outs.append(it.value().data.clone())
        it.close()
    assert torch.equal(outs[0], outs[1])


def test_input_dtype_bf16_batches_through_the_thread_buffer():
    """`input_dtype = bf16`: batches leave the chain as bf16 tensors holding the rounded fp32 values
    (half the host-to-device bytes); labels and indices are untouched"""
    want = [(b.data.clone(), b.label.clone(), b.inst_index.copy()) for b in _mk(SYN % "")]
    cfg = (SYN % "  input_dtype = bf16").replace("iter = end", "iter = threadbuffer\niter = end")
    it = create_iterator(parse_config_string(cfg))
    it.init()
    got = [(b.data.clone(), b.label.clone(), b.inst_index.copy()) for b in it]
    it.close()
    assert len(got) == len(want)
    for (d, l, i), (wd, wl, wi) in zip(got, want):
        assert d.dtype == torch.bfloat16 and torch.equal(d, wd.to(torch.bfloat16))
        assert torch.equal(l, wl) and (i == wi).all()
