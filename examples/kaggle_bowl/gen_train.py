#!/usr/bin/env python
"""usage: gen_train.py input_folder/ output_folder/ [size=48]
Resizes the class-per-directory training tree (same command line as the reference's gen_train.py,
which shells out to ImageMagick; the work is done by gen_resized.py with Pillow)."""
import sys

from gen_resized import main

if __name__ == "__main__":
    sys.exit(main(sys.argv))
