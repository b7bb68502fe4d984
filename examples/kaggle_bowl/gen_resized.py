#!/usr/bin/env python
"""usage: gen_resized.py input_folder/ output_folder/ [size=48]
Resize every image to size x size (the reference's gen_train.py / gen_test.py shell out to
ImageMagick ``convert -resize 48x48!``; this uses Pillow).  Works for class-per-directory trees
(train) and flat folders (test)."""
import os
import sys

from PIL import Image


def main(argv):
    if len(argv) < 3:
        print("Usage: python gen_resized.py input_folder output_folder [size]")
        return 1
    fi, fo = argv[1], argv[2]
    size = int(argv[3]) if len(argv) > 3 else 48
    for root, _, files in os.walk(fi):
        rel = os.path.relpath(root, fi)
        os.makedirs(os.path.join(fo, rel), exist_ok=True)
        for name in files:
            try:
                im = Image.open(os.path.join(root, name)).convert("RGB").resize((size, size))
                im.save(os.path.join(fo, rel, name))
            except OSError:
                pass
    return 0


if __name__ == "__main__":
    sys.exit(main(sys.argv))
