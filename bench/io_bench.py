#!/usr/bin/env python
"""Host data-pipeline throughput: imgbin pages of 256x256 JPEGs -> augmented 3x227x227 batches.

  python bench/io_bench.py [--n 2048] [--batch 256] [--threads 0]

Compares the Python iterator stages (`native_pipeline = 0`) with the fused native batch path
(csrc/runtime/batch_pipeline.cpp: JPEG decode + crop / mean / jitter / mirror / scale + batch assembly on worker
threads), fp32 and bf16 batches, and pages of pre-decoded raw records (`im2bin ... raw=1`).  CPU only; prints images/s."""
import argparse
import io
import os
import sys
import tempfile
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=2048)
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--threads", type=int, default=0)
    args = ap.parse_args()
    from PIL import Image
    from cxxnet_b200.config import parse_config_string
    from cxxnet_b200.io.data import create_iterator
    from cxxnet_b200.io.iterators import BinaryPage
    tmp = tempfile.mkdtemp()
    rs = np.random.RandomState(0)
    yy, xx = np.mgrid[0:256, 0:256]
    page, lines = BinaryPage(), []
    for i in range(args.n):
        a = np.stack([(np.sin(xx / (5.0 + i % 7)) + np.cos(yy / 9.0)) * 60 + 128, xx * 1.0, yy * 1.0], -1)
        a = np.clip(a + rs.randn(256, 256, 3) * 10, 0, 255).astype(np.uint8)
        b = io.BytesIO()
        Image.fromarray(a).save(b, "JPEG", quality=88)
        assert page.push(b.getvalue())
        lines.append("%d\t%d\tx\n" % (i, i % 1000))
    open(os.path.join(tmp, "a.lst"), "w").write("".join(lines))
    open(os.path.join(tmp, "a.bin"), "wb").write(bytes(page.raw))
    # the same images as pre-decoded raw records (what `im2bin ... raw=1` writes): several pages
    from cxxnet_b200.io import native
    with open(os.path.join(tmp, "raw.bin"), "wb") as f:
        rp = BinaryPage()
        for i in range(args.n):
            rec = native.raw_encode(np.asarray(Image.open(io.BytesIO(page.get(i))).convert("RGB")))
            if not rp.push(rec):
                f.write(bytes(rp.raw))
                rp = BinaryPage()
                assert rp.push(rec)
        f.write(bytes(rp.raw))
    print("cores: %d" % (os.cpu_count() or 1))
    for binfile, native_pipeline, dtype in (("a.bin", 0, "fp32"), ("a.bin", 1, "fp32"), ("a.bin", 1, "bf16"),
                                           ("raw.bin", 1, "bf16")):
        it = create_iterator(parse_config_string(
            'iter = imgbin\n image_list = "%s/a.lst"\n image_bin = "%s/%s"\n batch_size = %d\n'
            ' input_shape = 3,227,227\n rand_crop = 1\n rand_mirror = 1\n mean_value = 104,117,123\n silent = 1\n'
            ' native_pipeline = %d\n decode_threads = %d\n input_dtype = %s\niter = end\n'
            % (tmp, tmp, binfile, args.batch, native_pipeline, args.threads, dtype)))
        it.init()
        it.before_first()
        n = 0
        t0 = time.perf_counter()
        while it.next():
            n += it.value().batch_size
        dt = time.perf_counter() - t0
        it.close()
        print("%-8s native_pipeline=%d %s: %7.0f images/s (%d images)" % (binfile, native_pipeline, dtype, n / dt, n))


if __name__ == "__main__":
    main()
