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
