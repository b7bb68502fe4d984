#!/usr/bin/env python
"""conv1 of AlexNet as the space-to-depth stride-1 layer ([N][57][57][64] -> 55 x 55 x 96, 3 x 3, resident filter panel):
ring depth and rows per tile (cxn_conv_tune keys 0 and 1).  The layer is bound by load latency x bytes in flight."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

from cxxnet_b200 import ops
from gemm_sweep import timeit

L = ops.lib()
B = int(os.environ.get("BATCH", "256"))


def main():
    st = torch.cuda.current_stream().cuda_stream
    x = torch.randn(B, 57, 57, 64, device="cuda").to(torch.bfloat16)
    w = (torch.randn(96, 576, device="cuda") * 0.05).to(torch.bfloat16)
    bias = torch.zeros(96, device="cuda")
    y = torch.empty(B * 55 * 55, 96, device="cuda", dtype=torch.bfloat16)
    for mh, stages in [(1, 2), (1, 3), (1, 4), (1, 5), (1, 6), (1, 5), (1, 4)]:
        L.cxn_conv_tune(0, mh)
        L.cxn_conv_tune(1, stages)
        rc = L.cxn_conv_fprop(x.data_ptr(), B, 57, 57, 64, 64, 1, 8, w.data_ptr(), 576, 96, 3, 3, 1, 0, 0, 55, 55, y.data_ptr(), 96,
                              bias.data_ptr(), 1, st)
        if rc:
            print("rows/tile %s stages %s: rc %d" % (128 * mh or "auto", stages or "auto", rc))
            continue
        t = timeit(lambda: L.cxn_conv_fprop(x.data_ptr(), B, 57, 57, 64, 64, 1, 8, w.data_ptr(), 576, 96, 3, 3, 1, 0, 0, 55, 55,
                                            y.data_ptr(), 96, bias.data_ptr(), 1, st), iters=10)
        print("rows/tile %-4s stages %-4s: %.1f us" % (128 * mh or "auto", stages or "auto", t))
    L.cxn_conv_tune(0, 0)
    L.cxn_conv_tune(1, 0)


if __name__ == "__main__":
    main()
