#!/usr/bin/env python
"""usage: gen_test.py input_folder/ output_folder/ [size=48]
Resizes the flat folder of test images (same command line as the reference's gen_test.py; the work is
done by gen_resized.py with Pillow)."""
import sys

from gen_resized import main

if __name__ == "__main__":
    sys.exit(main(sys.argv))
