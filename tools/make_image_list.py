#!/usr/bin/env python
"""Image list (``index<TAB>label<TAB>relative/path``) from a class-per-directory tree, optionally resizing the images
on the way - the two preparation steps of examples/ImageNet/README.md.

  python tools/make_image_list.py ROOT out.lst [--val-fraction 0.02] [--shuffle 1] [--resize 256 --resized-root DIR]

* classes are the sub-directories of ROOT in sorted order (label = position); ``--synsets FILE`` writes that order
* ``--val-fraction f`` splits off every instance whose hash falls below f into ``<out>.val.lst`` (stable across runs)
* ``--resize N`` writes N x N JPEG copies (shorter side scaled to N, centre crop) under ``--resized-root`` and lists
  those; pack the result with ``bin/im2bin list resized-root out.bin [raw=1]``
"""
import argparse
import hashlib
import os
import random
import sys

EXT = (".jpg", ".jpeg", ".png", ".bmp")


def resize_to(src, dst, n):
    from PIL import Image
    im = Image.open(src).convert("RGB")
    w, h = im.size
    s = n / min(w, h)
    im = im.resize((max(n, int(round(w * s))), max(n, int(round(h * s)))), Image.BILINEAR)
    w, h = im.size
    x0, y0 = (w - n) // 2, (h - n) // 2
    os.makedirs(os.path.dirname(dst), exist_ok=True)
    im.crop((x0, y0, x0 + n, y0 + n)).save(dst, "JPEG", quality=92)


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("root")
    ap.add_argument("out")
    ap.add_argument("--val-fraction", type=float, default=0.0)
    ap.add_argument("--shuffle", type=int, default=1)
    ap.add_argument("--seed", type=int, default=888)
    ap.add_argument("--synsets", default="")
    ap.add_argument("--resize", type=int, default=0)
    ap.add_argument("--resized-root", default="")
    args = ap.parse_args(argv)
    if args.resize and not args.resized_root:
        ap.error("--resize needs --resized-root")
    classes = sorted(d for d in os.listdir(args.root) if os.path.isdir(os.path.join(args.root, d)))
    if not classes:
        print("no class directories under %s" % args.root, file=sys.stderr)
        return 1
    if args.synsets:
        with open(args.synsets, "w") as f:
            f.writelines(c + "\n" for c in classes)
    rows = []
    for label, cls in enumerate(classes):
        for name in sorted(os.listdir(os.path.join(args.root, cls))):
            if not name.lower().endswith(EXT):
                continue
            rel = cls + "/" + name
            if args.resize:
                out_rel = os.path.splitext(rel)[0] + ".jpg"
                try:
                    resize_to(os.path.join(args.root, rel), os.path.join(args.resized_root, out_rel), args.resize)
                except OSError as e:
                    print("skipping %s: %s" % (rel, e), file=sys.stderr)
                    continue
                rel = out_rel
            rows.append((label, rel))
    if args.shuffle:
        random.Random(args.seed).shuffle(rows)
    train, val = [], []
    for label, rel in rows:
        h = int(hashlib.md5(rel.encode()).hexdigest()[:8], 16) / float(1 << 32)
        (val if h < args.val_fraction else train).append((label, rel))
    stem = args.out[:-4] if args.out.endswith(".lst") else args.out
    outputs = [(args.out, train)] + ([(stem + ".val.lst", val)] if args.val_fraction else [])
    for path, part in outputs:
        with open(path, "w") as f:
            for i, (label, rel) in enumerate(part):
                f.write("%d\t%d\t%s\n" % (i, label, rel))
    print("%d classes, %d training and %d validation images" % (len(classes), len(train), len(val)))
    return 0


if __name__ == "__main__":
    sys.exit(main())
