#!/usr/bin/env python
"""What bounds the convolution weight-gradient kernel?  The same launch with parts switched off (cxn_conv_tune(5, dbg):
2 = the TMA loads are skipped, 4 = the MMAs are skipped; results are garbage, times are not): loads alone, MMAs +
epilogue alone, epilogue alone - for AlexNet's conv2..conv5 at batch BATCH, and with explicit split counts."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

from cxxnet_b200 import ops
from gemm_sweep import timeit

L = ops.lib()
B = int(os.environ.get("BATCH", "256"))
SHAPES = [("conv3", 13, 256, 1, 384, 3, 1), ("conv4", 13, 384, 2, 384, 3, 1), ("conv5", 13, 384, 2, 256, 3, 1),
          ("conv2", 27, 96, 2, 256, 5, 2)]


def main():
    st = torch.cuda.current_stream().cuda_stream
    for name, H, C, G, OC, k, pad in SHAPES:
        cg, ocg = C // G, OC // G
        x = torch.randn(B, H, H, C, device="cuda").to(torch.bfloat16)
        dy = torch.randn(B * H * H, OC, device="cuda").to(torch.bfloat16)
        kg = k * k * cg
        dw = torch.zeros(G, ocg, kg, device="cuda")
        for splits in (0, 5, 10, 40):
            out = []
            for dbg in (0, 2, 4, 6):
                L.cxn_conv_tune(5, dbg)
                t = timeit(lambda: ops.check(L.cxn_conv_wgrad(x.data_ptr(), B, H, H, C, cg, G, 8, dy.data_ptr(), OC, ocg, k, k, 1,
                                                              pad, pad, H, H, dw.data_ptr(), kg, splits, st), "wgrad"), iters=8)
                out.append(t)
            L.cxn_conv_tune(5, 0)
            print("%s wgrad b%d splits %2d: all %.1f us | no loads (MMA + epilogue) %.1f | no MMAs (loads + epilogue) %.1f | "
                  "epilogue alone %.1f" % (name, B, splits, out[0], out[1], out[2], out[3]))


if __name__ == "__main__":
    main()
