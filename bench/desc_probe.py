#!/usr/bin/env python
"""Does tcgen05.mma accept a 128B-swizzled K-major operand whose start address is shifted by whole 128-byte rows
(not aligned to the 1024-byte swizzle pattern), and does it need the descriptor's base-offset field?

  python bench/desc_probe.py          (on a B200)

For every shift 0..16 and both base-offset modes, prints whether the MMA read rows [shift, shift + 128) of the
patch.  The answer decides how the planned halo-tile convolution addresses its taps."""
import ctypes
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cxxnet_b200 import ops  # noqa: E402


def main():
    L = ops.lib()
    L.cxn_probe_shifted_desc.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
    L.cxn_probe_shifted_desc.restype = ctypes.c_int
    L.cxn_probe_expected.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int]
    L.cxn_probe_expected.restype = ctypes.c_float
    out = torch.zeros((128, 16), dtype=torch.float32, device="cuda")
    print("shift  base_offset=0  base_offset=(addr>>7)&7   (ok = rows [shift, shift+128) were read; else mismatching elements)")
    for shift in range(0, 17):
        cells = []
        for mode in (0, 1):
            out.zero_()
            ops.check(L.cxn_probe_shifted_desc(out.data_ptr(), shift, mode, ops.stream()), "probe")
            torch.cuda.synchronize()
            want = torch.tensor([[L.cxn_probe_expected(m, n, shift) for n in range(16)] for m in range(128)])
            bad = int((out.cpu() != want).sum())
            cells.append("ok" if bad == 0 else "%d/2048 differ" % bad)
        print("%5d  %-13s  %s" % (shift, cells[0], cells[1]))


if __name__ == "__main__":
    main()
