#!/usr/bin/env python
"""split-K sweep of the convolution weight gradient on the AlexNet shapes"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

from cxxnet_b200 import ops

L = ops.lib()
B = int(os.environ.get("BATCH", "256"))
SHAPES = [("conv2", 27, 96, 2, 256, 5, 2), ("conv3", 13, 256, 1, 384, 3, 1), ("conv4", 13, 384, 2, 384, 3, 1),
          ("conv5", 13, 384, 2, 256, 3, 1)]


def timeit(run, iters=10):
    for _ in range(3):
        run()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        run()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / iters


st = torch.cuda.current_stream().cuda_stream
for name, H, C, G, OC, k, pad in SHAPES:
    cg, ocg = C // G, OC // G
    kg = k * k * cg
    x = torch.randn(B, H, H, C, device="cuda").to(torch.bfloat16)
    dy = torch.randn(B * H * H, OC, device="cuda").to(torch.bfloat16)
    dw = torch.zeros(G, ocg, kg, device="cuda")
    flops = 2.0 * B * H * H * OC * kg
    line = [name]
    for sp in (0, 4, 6, 8, 10, 12, 16, 21, 32, 42):
        us = timeit(lambda: L.cxn_conv_wgrad(x.data_ptr(), B, H, H, C, cg, G, 8, dy.data_ptr(), OC, ocg, k, k, 1, pad, pad, H, H,
                                             dw.data_ptr(), kg, sp, st))
        line.append("%s:%.0fus" % ("auto" if sp == 0 else sp, us))
    print(" ".join(line), "(%.1f GF)" % (flops / 1e9), flush=True)
