"""Native JPEG decoder (csrc/runtime/jpeg_decoder.cpp, the counterpart of the reference's libjpeg-backed
utils/decoder.h) against Pillow's libjpeg on baseline / progressive, 4:4:4 / 4:2:2 / 4:2:0, grey-scale,
restart intervals and sizes that are not multiples of the MCU; the threaded batch entry point; the fallback
for inputs it does not cover; and imgbin / imgbinx iterators over JPEG pages."""
import io

import numpy as np
import pytest

PIL = pytest.importorskip("PIL")
from PIL import Image  # noqa: E402

from cxxnet_b200.io import native  # noqa: E402

pytestmark = pytest.mark.skipif(native.lib() is None or not hasattr(native.lib(), "cxnio_jpeg_info"),
                                reason="libcxxnet_io.so not built")


def _picture(w, h, seed=0, gray=False):
    rs = np.random.RandomState(seed)
    yy, xx = np.mgrid[0:h, 0:w]
    a = np.stack([(np.sin(xx / 7.0) + np.cos(yy / 5.0)) * 60 + 128, xx * 255.0 / max(w - 1, 1),
                  yy * 255.0 / max(h - 1, 1)], -1) + rs.randn(h, w, 3) * 12
    im = Image.fromarray(np.clip(a, 0, 255).astype(np.uint8))
    return im.convert("L") if gray else im


def _encode(im, **kw):
    b = io.BytesIO()
    im.save(b, "JPEG", **kw)
    return b.getvalue()


def _reference(buf):
    return np.asarray(Image.open(io.BytesIO(buf)).convert("RGB")).astype(np.int32).transpose(2, 0, 1)


@pytest.mark.parametrize("size", [(64, 48), (67, 45), (1, 1), (17, 9), (255, 129), (8, 8)])
@pytest.mark.parametrize("progressive", [False, True])
def test_matches_libjpeg(size, progressive):
    w, h = size
    for sub in (0, 1, 2):
        for q in (50, 92):
            for rst in (0, 2):
                kw = dict(quality=q, progressive=progressive, subsampling=sub)
                if rst:
                    kw["restart_marker_blocks"] = rst
                buf = _encode(_picture(w, h, seed=sub * 7 + q), **kw)
                assert native.jpeg_info(buf) == (w, h, 3)
                got = native.jpeg_decode(buf)
                assert got is not None and got.dtype == np.float32 and got.shape == (3, h, w)
                d = np.abs(got.astype(np.int32) - _reference(buf))
                # same entropy decoding and up-sampling; the inverse DCT is a float AAN transform instead of
                # libjpeg's fixed-point one, which the colour conversion can amplify to 3 levels
                assert d.max() <= 3 and d.mean() < 0.25, (size, progressive, sub, q, rst, d.max(), d.mean())


def test_grayscale_and_batch_threads():
    bufs = [_encode(_picture(40 + i, 30 + 2 * i, seed=i, gray=(i % 3 == 0)), quality=85, progressive=(i % 2 == 1))
            for i in range(13)]
    outs = native.jpeg_decode_many(bufs, nthreads=4)
    assert len(outs) == 13
    for i, (b, o) in enumerate(zip(bufs, outs)):
        assert o is not None and o.shape == (3, 30 + 2 * i, 40 + i)
        assert np.abs(o.astype(np.int32) - _reference(b)).max() <= 3
        assert np.array_equal(o, native.jpeg_decode(b))          # threads do not change results


def test_unsupported_inputs_fall_back():
    from cxxnet_b200.io.iterators import _decode_image
    png = io.BytesIO()
    _picture(20, 10).save(png, "PNG")
    assert native.jpeg_decode(png.getvalue()) is None and native.jpeg_info(b"\xff\xd8\xff") is None
    assert native.image_decode(png.getvalue()).shape == (3, 10, 20)      # PNG has its own native decoder
    cmyk = _encode(_picture(16, 16).convert("CMYK"), quality=90)
    assert native.jpeg_decode(cmyk) is None                      # four components: left to the library decoders
    truncated = _encode(_picture(64, 64), quality=90)[:200]
    assert native.jpeg_decode(truncated) is None or native.jpeg_decode(truncated).shape == (3, 64, 64)
    mixed = native.jpeg_decode_many([png.getvalue(), _encode(_picture(9, 9), quality=80)])
    assert mixed[0] is None and mixed[1].shape == (3, 9, 9)
    arr = _decode_image(png.getvalue())                          # PNG still decodes through OpenCV / Pillow
    assert arr.shape == (3, 10, 20) and arr.dtype == np.float32


def test_imgbin_iterators_decode_jpeg_pages(tmp_path):
    from cxxnet_b200.config import parse_config_string
    from cxxnet_b200.io.data import create_iterator
    from cxxnet_b200.io.iterators import BinaryPage
    page, lines, bufs = BinaryPage(), [], []
    for i in range(7):
        bufs.append(_encode(_picture(24, 20, seed=i), quality=90, subsampling=2))
        assert page.push(bufs[-1])
        lines.append("%d\t%d\timg%d.jpg\n" % (i, i % 4, i))
    (tmp_path / "a.lst").write_text("".join(lines))
    (tmp_path / "a.bin").write_bytes(bytes(page.raw))
    for kind in ("imgbin", "imgbinx"):
        it = create_iterator(parse_config_string(
            'iter = %s\n image_list = "%s"\n image_bin = "%s"\n batch_size = 3\n input_shape = 3,20,24\n'
            ' round_batch = 0\n silent = 1\niter = end\n' % (kind, tmp_path / "a.lst", tmp_path / "a.bin")))
        it.init()
        it.before_first()
        rows = []
        while it.next():
            b = it.value()
            for r in range(b.batch_size - b.num_batch_padd):
                rows.append((int(b.inst_index[r]), b.data[r].numpy().copy()))
        it.close()
        assert [r[0] for r in rows[:6]] == [0, 1, 2, 3, 4, 5]
        for idx, data in rows[:6]:
            assert np.abs(data.astype(np.int32) - _reference(bufs[idx])).max() <= 3


def _jpeg_bin(tmp_path, n=11, w=30, h=28):
    from cxxnet_b200.io.iterators import BinaryPage
    page, lines = BinaryPage(), []
    for i in range(n):
        if i == 4:                               # one PNG among the JPEGs: decoded by the library path
            b = io.BytesIO()
            _picture(w, h, seed=i).save(b, "PNG")
            blob = b.getvalue()
        else:
            blob = _encode(_picture(w, h, seed=i), quality=88, subsampling=i % 3, progressive=(i % 2 == 0))
        assert page.push(blob)
        lines.append("%d\t%d\timg%d\n" % (i, i % 5, i))
    (tmp_path / "j.lst").write_text("".join(lines))
    (tmp_path / "j.bin").write_bytes(bytes(page.raw))


@pytest.mark.parametrize("mean", ["none", "value", "image"])
@pytest.mark.parametrize("dtype", ["fp32", "bf16"])
def test_native_batch_pipeline_equals_python_stages(tmp_path, mean, dtype):
    """decode + augment + batch assembly on native threads (csrc/runtime/batch_pipeline.cpp) produces the batches
    of the Python iterator stages bit for bit: same random decisions, same single-precision arithmetic"""
    import torch
    from cxxnet_b200.config import parse_config_string
    from cxxnet_b200.io.data import create_iterator
    _jpeg_bin(tmp_path)
    extra = {"none": "", "value": " mean_value = 100,110,120\n",
             "image": ' image_mean = "%s"\n' % (tmp_path / "mean.bin")}[mean]

    def batches(native_pipeline, kind):
        it = create_iterator(parse_config_string(
            'iter = %s\n image_list = "%s"\n image_bin = "%s"\n batch_size = 4\n input_shape = 3,24,26\n'
            ' rand_crop = 1\n rand_mirror = 1\n max_random_contrast = 0.2\n max_random_illumination = 6\n'
            ' divideby = 64\n round_batch = 1\n seed_data = 5\n silent = 1\n input_dtype = %s\n native_pipeline = %d\n'
            ' decode_threads = 3\n%siter = end\n'
            % (kind, tmp_path / "j.lst", tmp_path / "j.bin", dtype, native_pipeline, extra)))
        it.init()
        assert it._native == bool(native_pipeline)
        out = []
        for _ in range(2):                       # two passes: the wrap-around of round_batch included
            it.before_first()
            while it.next():
                b = it.value()
                out.append((b.data.clone(), b.label.clone(), b.inst_index.copy(), b.num_batch_padd))
        it.close()
        return out

    if mean == "image":
        batches(0, "imgbin")                     # the first use computes and caches the mean image (and draws while doing so)
        assert (tmp_path / "mean.bin").exists()
    want = batches(0, "imgbin")
    for kind in ("imgbin", "imgbinx"):
        got = batches(1, kind)
        assert len(got) == len(want) == 6
        for (d0, l0, i0, p0), (d1, l1, i1, p1) in zip(want, got):
            assert d1.dtype == (torch.bfloat16 if dtype == "bf16" else torch.float32)
            assert torch.equal(d0, d1) and torch.equal(l0, l1) and list(i0) == list(i1) and p0 == p1


def test_native_augment_modes_against_numpy_statement():
    """every mean mode of cxnio_decode_augment_batch on decoded pixels and on JPEG bytes, mirrored or not,
    against AugmentIterator.apply_params"""
    import torch
    from cxxnet_b200.io.iterators import AugmentIterator, SyntheticInstIterator
    rs = np.random.RandomState(3)
    H, W = 20, 22
    for mode_name in ("none", "value", "mean_full", "mean_crop"):
        aug = AugmentIterator(SyntheticInstIterator())
        aug.shape = (3, H, W)
        aug.scale = 1.0 / 32
        if mode_name == "value":
            aug.mean_b, aug.mean_g, aug.mean_r = 90.0, 100.0, 110.0
        if mode_name.startswith("mean_"):
            aug.name_meanimg, aug.meanfile_ready = "x", True
            aug.meanimg = (rs.rand(3, 26, 30) * 255).astype(np.float32) if mode_name == "mean_full" \
                else (rs.rand(3, H, W) * 255).astype(np.float32)
        items, want = [], []
        for i in range(6):
            blob = _encode(_picture(30, 26, seed=20 + i), quality=90)
            pixels = native.jpeg_decode(blob)
            params = (i % 5, (2 * i) % 7, i % 2 == 1, 1.0 + 0.03 * i, 2.0 - i)
            items.append((blob, pixels if i % 3 == 0 else None, 26, 30) + params)
            want.append(aug.apply_params(pixels, params))
        mode = aug.mean_mode()
        mean = {0: None, 1: np.asarray([90.0, 100.0, 110.0], np.float32), 2: aug.meanimg}[mode]
        for bf16 in (False, True):
            dst = torch.zeros((6, 3, H, W), dtype=torch.bfloat16 if bf16 else torch.float32)
            st = native.decode_augment_batch(items, (3, H, W), mode, mean, aug.scale, dst, bf16, 2)
            assert st == [0] * 6
            ref = torch.from_numpy(np.stack(want))
            assert torch.equal(dst, ref.to(dst.dtype)), mode_name
    # a crop that leaves the image is reported, not written
    dst = torch.zeros((1, 3, H, W))
    bad = [(blob, None, 26, 30, 10, 0, False, 1.0, 0.0)]
    assert native.decode_augment_batch(bad, (3, H, W), 0, None, 1.0, dst, False, 1) != [0]


def test_decoder_survives_mutated_inputs_under_sanitizers(tmp_path):
    """tools/jpeg_fuzz.cpp built with ASAN + UBSAN: damaged files may be rejected, never crash or over-read"""
    import os
    import shutil
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cxx = shutil.which("g++")
    if cxx is None:
        pytest.skip("no g++")
    exe = tmp_path / "jpeg_fuzz"
    r = subprocess.run([cxx, "-O1", "-g", "-std=c++17", "-fsanitize=address,undefined",
                        "-fno-sanitize-recover=undefined", "-pthread", "-o", str(exe),
                        os.path.join(root, "tools", "jpeg_fuzz.cpp"),
                        os.path.join(root, "csrc", "runtime", "jpeg_decoder.cpp"),
                        os.path.join(root, "csrc", "runtime", "png_decoder.cpp"), "-lz"],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
    if r.returncode != 0 and "sanitize" in r.stdout:
        pytest.skip("sanitizer runtime not available: " + r.stdout[-200:])
    assert r.returncode == 0, r.stdout[-2000:]
    seeds = []
    for i, kw in enumerate([dict(quality=90), dict(quality=80, progressive=True),
                            dict(quality=70, subsampling=2, restart_marker_blocks=2),
                            dict(quality=95, subsampling=1, progressive=True)]):
        p = tmp_path / ("seed%d.jpg" % i)
        p.write_bytes(_encode(_picture(52, 40, seed=i), **kw))
        seeds.append(str(p))
    for i, mode in enumerate(["RGB", "P", "LA", "1"]):
        p = tmp_path / ("seed%d.png" % i)
        im = _picture(52, 40, seed=i)
        (im.quantize(colors=17) if mode == "P" else im.convert(mode)).save(p, "PNG")
        seeds.append(str(p))
    r = subprocess.run([str(exe), "1500"] + seeds, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True,
                       timeout=600)
    assert r.returncode == 0 and "decoded" in r.stdout, r.stdout[-3000:]
    decoded = int(r.stdout.split("decoded")[1].split()[0])
    assert decoded > 100          # the mutations are mild enough that a good share still decodes


def test_rank_slice_decodes_only_the_rows_a_rank_uses(tmp_path):
    """slice_world / slice_rank (set by the task driver in multi-process runs): the rows of a rank's slice are those
    of the unsliced batch, the random decisions do not depend on the split, the other rows are not decoded"""
    import torch
    from cxxnet_b200.config import parse_config_string
    from cxxnet_b200.io.data import create_iterator
    _jpeg_bin(tmp_path, n=12)

    def batches(extra):
        it = create_iterator(parse_config_string(
            'iter = imgbin\n image_list = "%s"\n image_bin = "%s"\n batch_size = 6\n input_shape = 3,24,26\n'
            ' rand_crop = 1\n rand_mirror = 1\n mean_value = 100,110,120\n seed_data = 9\n silent = 1\n%siter = end\n'
            % (tmp_path / "j.lst", tmp_path / "j.bin", extra)))
        it.init()
        it.before_first()
        out = []
        while it.next():
            b = it.value()
            out.append((b.data.clone(), b.label.clone(), b.inst_index.copy()))
            b.data.fill_(-777.0)                  # marks the rows the next batch leaves alone
        it.close()
        return out

    full = batches("")
    assert len(full) == 2
    for rank in range(3):                         # 3 ranks: rows [0,2) [2,4) [4,6)
        part = batches(" slice_world = 3\n slice_rank = %d\n" % rank)
        for (d0, l0, i0), (d1, l1, i1) in zip(full, part):
            lo, hi = 2 * rank, 2 * rank + 2
            assert torch.equal(d0[lo:hi], d1[lo:hi]) and torch.equal(l0, l1) and list(i0) == list(i1)
        other = [r for r in range(6) if not 2 * rank <= r < 2 * rank + 2]
        assert bool((part[1][0][other] == -777.0).all())       # second batch: untouched rows still hold the marker


def test_native_path_keeps_the_order_of_a_shuffled_source_behind_a_thread_buffer(tmp_path):
    """imgbinx with shuffle = 1 wrapped in threadbuffer: the native batch path visits the data in the same
    (shuffled) order and produces the same batches as the Python stages, over two passes"""
    import torch
    from cxxnet_b200.config import parse_config_string
    from cxxnet_b200.io.data import create_iterator
    _jpeg_bin(tmp_path, n=13)

    def run(native_pipeline):
        it = create_iterator(parse_config_string(
            'iter = imgbinx\n image_list = "%s"\n image_bin = "%s"\n shuffle = 1\n batch_size = 4\n'
            ' input_shape = 3,24,26\n rand_crop = 1\n rand_mirror = 1\n mean_value = 1,2,3\n round_batch = 1\n'
            ' silent = 1\n native_pipeline = %d\niter = threadbuffer\n silent = 1\niter = end\n'
            % (tmp_path / "j.lst", tmp_path / "j.bin", native_pipeline)))
        it.init()
        out = []
        for _ in range(2):
            it.before_first()
            while it.next():
                b = it.value()
                out.append((b.data.clone(), [int(i) for i in b.inst_index], b.num_batch_padd))
        it.close()
        return out

    a, b = run(0), run(1)
    assert len(a) == len(b) > 4
    assert [x[1] for x in a] != [sorted(x[1]) for x in a]        # the order really is shuffled
    for (d0, i0, p0), (d1, i1, p1) in zip(a, b):
        assert i0 == i1 and p0 == p1 and torch.equal(d0, d1)


def test_command_line_trains_from_packed_jpegs(tmp_path):
    """files -> bin/im2bin -> `iter = imgbin` (native reader, decoder and batch path) -> task driver on the CPU
    backend: a two-class problem (bright / dark images) is learnt, and `task = pred` reads the same pages"""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    im2bin = os.path.join(root, "bin", "im2bin")
    if not os.path.exists(im2bin):
        pytest.skip("bin/im2bin not built")
    rs = np.random.RandomState(0)
    os.makedirs(tmp_path / "img")
    lines = []
    for i in range(60):
        cls = i % 2
        a = np.clip(rs.randn(20, 22, 3) * 20 + (70 if cls == 0 else 180), 0, 255).astype(np.uint8)
        Image.fromarray(a).save(tmp_path / "img" / ("%03d.jpg" % i), quality=90)
        lines.append("%d\t%d\timg/%03d.jpg\n" % (i, cls, i))
    (tmp_path / "all.lst").write_text("".join(lines))
    subprocess.run([im2bin, str(tmp_path / "all.lst"), str(tmp_path) + "/", str(tmp_path / "all.bin")], check=True,
                   stdout=subprocess.PIPE, timeout=120)
    conf = tmp_path / "net.conf"
    conf.write_text('''
data = train
iter = imgbin
  image_list = "%(d)s/all.lst"
  image_bin = "%(d)s/all.bin"
  rand_crop = 1
  rand_mirror = 1
  mean_value = 128,128,128
  divideby = 64
iter = threadbuffer
iter = end
eval = val
iter = imgbin
  image_list = "%(d)s/all.lst"
  image_bin = "%(d)s/all.bin"
  mean_value = 128,128,128
  divideby = 64
iter = end
pred = %(d)s/pred.txt
iter = imgbin
  image_list = "%(d)s/all.lst"
  image_bin = "%(d)s/all.bin"
  mean_value = 128,128,128
  divideby = 64
iter = end
netconfig = start
layer[+1] = flatten
layer[+1] = fullc:fc
  nhidden = 2
  init_sigma = 0.01
layer[+0] = softmax
netconfig = end
input_shape = 3,16,16
batch_size = 20
round_batch = 1
dev = cpu
eta = 0.05
momentum = 0.9
metric = error
num_round = 3
max_round = 3
model_dir = %(d)s/models
silent = 1
''' % {"d": tmp_path})
    env = dict(os.environ)
    env["PYTHONPATH"] = root + os.pathsep + env.get("PYTHONPATH", "")
    r = subprocess.run([sys.executable, "-m", "cxxnet_b200", str(conf)], cwd=tmp_path, env=env, stdout=subprocess.PIPE,
                       stderr=subprocess.PIPE, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    last = [ln for ln in r.stderr.splitlines() if ln.startswith("[3]")][0]
    assert float(last.split("val-error:")[1].split()[0]) == 0.0, last
    r = subprocess.run([sys.executable, "-m", "cxxnet_b200", str(conf), "task=pred",
                        "model_in=%s/models/0003.model" % tmp_path], cwd=tmp_path, env=env, stdout=subprocess.PIPE,
                       stderr=subprocess.PIPE, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    preds = [int(float(x)) for x in (tmp_path / "pred.txt").read_text().split()]
    assert preds == [i % 2 for i in range(60)]


def test_native_affine_stage_equals_the_numpy_warp(tmp_path):
    """rotation / shear / aspect / scale (the kaggle bowl augmentation) on the native batch path: same random
    decisions, same bicubic (default) / bilinear arithmetic, bit-identical batches"""
    import torch
    from cxxnet_b200.config import parse_config_string
    from cxxnet_b200.io.data import create_iterator
    _jpeg_bin(tmp_path, n=10, w=48, h=48)

    def batches(native_pipeline, extra):
        it = create_iterator(parse_config_string(
            'iter = imgbin\n image_list = "%s"\n image_bin = "%s"\n batch_size = 5\n input_shape = 3,40,40\n'
            ' rand_crop = 1\n rand_mirror = 1\n max_rotate_angle = 180\n max_aspect_ratio = 0.5\n'
            ' max_shear_ratio = 0.3\n min_crop_size = 32\n max_crop_size = 48\n seed_data = 4\n silent = 1\n'
            ' native_pipeline = %d\n%siter = end\n' % (tmp_path / "j.lst", tmp_path / "j.bin", native_pipeline, extra)))
        it.init()
        assert it._native == bool(native_pipeline)
        out = []
        it.before_first()
        while it.next():
            out.append(it.value().data.clone())
        it.close()
        return out

    for extra in ("", " mean_value = 90,100,110\n divideby = 32\n max_random_contrast = 0.1\n",
                  ' image_mean = "%s"\n fill_value = 0\n' % (tmp_path / "m.bin"),
                  " warp_interp = bilinear\n"):      # default: bicubic (cv::INTER_CUBIC of the reference)
        if "image_mean" in extra:
            batches(0, extra)                       # first use creates the mean image
        a, b = batches(0, extra), batches(1, extra)
        assert len(a) == len(b) == 2
        for x, y in zip(a, b):
            assert torch.equal(x, y), float((x - y).abs().max())
    assert float(a[0].std()) > 1.0


def test_raw_records_skip_the_decoding(tmp_path):
    """`im2bin ... raw=1` decodes at packing time; the iterators (native batch path and Python stages) read the raw
    records and deliver exactly what they deliver from the JPEG pages"""
    import os
    import subprocess
    import torch
    from cxxnet_b200.config import parse_config_string
    from cxxnet_b200.io.data import create_iterator
    from cxxnet_b200.io.iterators import _decode_image
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    im2bin = os.path.join(root, "bin", "im2bin")
    if not os.path.exists(im2bin):
        pytest.skip("bin/im2bin not built")
    os.makedirs(tmp_path / "img")
    lines = []
    for i in range(9):
        (tmp_path / "img" / ("%d.jpg" % i)).write_bytes(_encode(_picture(30, 28, seed=i), quality=90, progressive=(i % 2 == 1)))
        lines.append("%d\t%d\timg/%d.jpg\n" % (i, i % 3, i))
    (tmp_path / "a.lst").write_text("".join(lines))
    for name, extra in (("jpg.bin", []), ("raw.bin", ["raw=1"])):
        subprocess.run([im2bin, str(tmp_path / "a.lst"), str(tmp_path) + "/", str(tmp_path / name)] + extra, check=True,
                       stdout=subprocess.PIPE, timeout=120)
    assert os.path.getsize(tmp_path / "raw.bin") == os.path.getsize(tmp_path / "jpg.bin") == 64 << 20

    def batches(binfile, native_pipeline):
        it = create_iterator(parse_config_string(
            'iter = imgbin\n image_list = "%s"\n image_bin = "%s"\n batch_size = 4\n input_shape = 3,24,26\n rand_crop = 1\n'
            ' rand_mirror = 1\n mean_value = 100,110,120\n round_batch = 1\n silent = 1\n native_pipeline = %d\niter = end\n'
            % (tmp_path / "a.lst", tmp_path / binfile, native_pipeline)))
        it.init()
        it.before_first()
        out = []
        while it.next():
            out.append(it.value().data.clone())
        it.close()
        return out
    want = batches("jpg.bin", 1)
    for native_pipeline in (0, 1):
        got = batches("raw.bin", native_pipeline)
        assert len(got) == len(want) == 3 and all(torch.equal(a, b) for a, b in zip(want, got))
    # record round trip through the library and the NumPy reader
    px = (np.random.RandomState(0).rand(5, 7, 3) * 255).astype(np.uint8)
    rec = native.raw_encode(px)
    assert rec[:4] == b"CXNR" and len(rec) == 12 + 5 * 7 * 3 and native.image_info(rec) == (7, 5, 3)
    want_planar = px.transpose(2, 0, 1).astype(np.float32)
    assert np.array_equal(native.image_decode(rec), want_planar)
    assert np.array_equal(_decode_image(rec, native_first=False), want_planar)
    assert native.image_info(rec[:-1]) is None                       # truncated record


@pytest.mark.parametrize("mode", ["RGB", "L", "P", "RGBA", "LA", "1", "I;16"])
def test_png_decoder_matches_pillow(mode):
    """native PNG path (csrc/runtime/png_decoder.cpp): colour types, palette, alpha dropped, packed bits, every
    scanline filter (the encoder picks them adaptively on noisy gradients)"""
    rs = np.random.RandomState(3)
    for (w, h) in [(31, 17), (64, 40), (5, 3), (1, 1)]:
        yy, xx = np.mgrid[0:h, 0:w]
        base = np.clip(np.stack([xx * 255.0 / max(w - 1, 1), yy * 255.0 / max(h - 1, 1), (xx + yy) * 4.0], -1)
                       + rs.randn(h, w, 3) * 6, 0, 255).astype(np.uint8)
        im = Image.fromarray(base)
        if mode == "I;16":
            im = Image.fromarray((base[:, :, 0].astype(np.uint16) << 8) + 37)
        elif mode == "P":
            im = im.quantize(colors=23)
        else:
            im = im.convert(mode)
        b = io.BytesIO()
        im.save(b, "PNG")
        buf = b.getvalue()
        assert native.image_info(buf) == (w, h, 3)
        got = native.image_decode(buf)
        if mode == "I;16":
            want = np.repeat(base[:, :, :1], 3, axis=2)              # high byte of every sample
        elif mode in ("RGBA", "LA"):
            want = np.asarray(im.convert("RGBA"))[:, :, :3]          # alpha dropped, not blended
        else:
            want = np.asarray(im.convert("RGB"))
        assert got is not None and np.array_equal(got, want.transpose(2, 0, 1).astype(np.float32)), (mode, w, h)


def test_png_interlaced_or_damaged_falls_back():
    from cxxnet_b200.io.iterators import _decode_image
    b = io.BytesIO()
    _picture(20, 10).save(b, "PNG")
    good = b.getvalue()
    cut = good[:len(good) // 2]
    assert native.image_decode(cut) is None
    flipped = bytearray(good)
    flipped[len(good) // 2] ^= 0x55
    out = native.image_decode(bytes(flipped))
    assert out is None or out.shape == (3, 10, 20)                   # zlib / filter damage is reported, never a crash
    assert np.array_equal(_decode_image(good), np.asarray(_picture(20, 10)).transpose(2, 0, 1).astype(np.float32))


def test_u8_batches_carry_raw_pixels_and_their_normalisation(tmp_path):
    """input_dtype = u8: cropped / mirrored 8-bit pixels plus DataBatch.norm; (pixel - mean) * scale reproduces the
    fp32 batch exactly, through a thread buffer too; configurations whose augmentation changes pixel values are
    refused; the CPU backend trains to the same weights from either kind of batch"""
    import torch
    from cxxnet_b200.config import parse_config_string
    from cxxnet_b200.io.data import create_iterator
    _jpeg_bin(tmp_path, n=10)
    base = ('iter = imgbin\n image_list = "%s"\n image_bin = "%s"\n batch_size = 4\n input_shape = 3,24,26\n rand_crop = 1\n'
            ' rand_mirror = 1\n mean_value = 100,110,120\n divideby = 32\n round_batch = 1\n seed_data = 2\n silent = 1\n'
            % (tmp_path / "j.lst", tmp_path / "j.bin"))

    def batches(extra):
        it = create_iterator(parse_config_string(base + extra + "iter = threadbuffer\n silent = 1\niter = end\n"))
        it.init()
        it.before_first()
        out = []
        while it.next():
            b = it.value()
            out.append((b.data.clone(), b.norm, b.label.clone()))
        it.close()
        return out
    f32, u8 = batches(""), batches(" input_dtype = u8\n")
    assert len(f32) == len(u8) == 3
    for (d0, n0, l0), (d1, n1, l1) in zip(f32, u8):
        assert n0 is None and d1.dtype == torch.uint8 and torch.equal(l0, l1)
        mean, scale = n1
        assert mean == [100.0, 110.0, 120.0] and scale == 1.0 / 32
        assert torch.equal((d1.float() - torch.tensor(mean).view(1, 3, 1, 1)) * scale, d0)
    with pytest.raises(ValueError, match="u8"):
        create_iterator(parse_config_string(base + " input_dtype = u8\n max_random_contrast = 0.2\niter = end\n")).init()

    # the reference-math backend trains identically from 8-bit and from float batches
    sys_path = __import__("os").path.join(__import__("os").path.dirname(__file__))
    __import__("sys").path.insert(0, sys_path)
    from util import make_trainer
    conf = """
netconfig=start
layer[+1] = flatten
layer[+1] = fullc:fc
  nhidden = 5
  init_sigma = 0.01
layer[+0] = softmax
netconfig=end
input_shape = 3,24,26
batch_size = 4
dev = cpu
eta = 0.01
metric = error
"""
    ta, tb = make_trainer(conf), make_trainer(conf)
    tb.set_weight(ta.get_weight("fc", "wmat")[0], "fc", "wmat")
    for kind, tr in (("", ta), (" input_dtype = u8\n", tb)):
        it = create_iterator(parse_config_string(base + kind + "iter = end\n"))
        it.init()
        it.before_first()
        while it.next():
            tr.update(it.value())
        it.close()
    assert np.array_equal(ta.get_weight("fc", "wmat")[0], tb.get_weight("fc", "wmat")[0])


def test_pipeline_threads_are_race_free_under_thread_sanitizer(tmp_path):
    """tools/io_tsan.cpp built with -fsanitize=thread: page reader thread (record and page interfaces, rewind in
    mid-page) and concurrent batch calls with their worker threads"""
    import os
    import shutil
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cxx = shutil.which("g++")
    if cxx is None:
        pytest.skip("no g++")
    exe = tmp_path / "io_tsan"
    srcs = [os.path.join(root, "tools", "io_tsan.cpp")] + [os.path.join(root, "csrc", "runtime", f) for f in
            ("page_reader.cpp", "jpeg_decoder.cpp", "png_decoder.cpp", "raw_image.cpp", "batch_pipeline.cpp")]
    r = subprocess.run([cxx, "-O1", "-g", "-std=c++17", "-fsanitize=thread", "-DCXNIO_NO_CLONES", "-pthread", "-o", str(exe)]
                       + srcs + ["-lz"], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
    if r.returncode != 0 and "sanitize" in r.stdout:
        pytest.skip("sanitizer runtime not available: " + r.stdout[-200:])
    assert r.returncode == 0, r.stdout[-2000:]
    r = subprocess.run([str(exe), str(tmp_path / "scratch.bin")], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True,
                       timeout=600)
    if r.returncode != 0 and "unexpected memory mapping" in r.stdout:
        pytest.skip("thread sanitizer cannot map its shadow memory on this kernel")
    assert r.returncode == 0 and r.stdout.strip().endswith("ok"), r.stdout[-3000:]


def test_avx2_idct_is_bit_identical_to_the_portable_one():
    """the AVX2 inverse DCT performs the portable code's operations in the same order: not one byte may differ over
    sparse, dense and out-of-range coefficient blocks (skipped on CPUs without AVX2)"""
    import ctypes
    from cxxnet_b200.io import native
    L = ctypes.CDLL(native._LIB_PATH) if hasattr(native, "_LIB_PATH") else None
    if L is None:
        import os
        L = ctypes.CDLL(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(native.__file__))), "lib", "libcxxnet_io.so"))
    L.cxnio_jpeg_idct_selftest.argtypes = [ctypes.c_int, ctypes.c_uint]
    L.cxnio_jpeg_idct_selftest.restype = ctypes.c_int
    for seed in (1, 7, 1234):
        r = L.cxnio_jpeg_idct_selftest(20000, seed)
        if r < 0:
            pytest.skip("no AVX2 on this CPU")
        assert r == 0
