#!/usr/bin/env python
"""A/B two builds of the kernel library on the AlexNet convolution problems (plain C ABI calls).
usage: conv_ab.py libA.so [libB.so ...]"""
import ctypes
import sys

import torch

P, I, LL = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong
B = 256
# name, H, C(pitch), Cg, G, OC, k, stride, pad, gran
LAYERS = [("conv1", 227, 4, 4, 1, 96, 11, 4, 0, 4), ("conv2", 27, 96, 48, 2, 256, 5, 1, 2, 8), ("conv2", 27, 96, 48, 2, 256, 5, 1, 2, 8),
          ("conv3", 13, 256, 256, 1, 384, 3, 1, 1, 8), ("conv4", 13, 384, 192, 2, 384, 3, 1, 1, 8),
          ("conv5", 13, 384, 192, 2, 256, 3, 1, 1, 8)]


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


def main():
    libs = []
    for path in sys.argv[1:]:
        L = ctypes.CDLL(path)
        L.cxn_conv_fprop.argtypes = [P, I, I, I, I, I, I, I, P, LL, I, I, I, I, I, I, I, I, P, LL, P, I, P]
        L.cxn_conv_wgrad.argtypes = [P, I, I, I, I, I, I, I, P, LL, I, I, I, I, I, I, I, I, P, LL, I, P]
        libs.append((path, L))
    st = torch.cuda.current_stream().cuda_stream
    for name, H, C, cg, G, OC, k, s, pad, gran in LAYERS:
        OH = (H + 2 * pad - k) // s + 1
        ocg = OC // G
        kg = (k * k * cg + 7) // 8 * 8      # leading dimension of the filter matrix (16-byte rows)
        x = torch.randn(B, H, H, C, device="cuda").to(torch.bfloat16)
        w = (torch.randn(G, ocg, kg, device="cuda") * 0.05).to(torch.bfloat16)
        y = torch.empty(B * OH * OH, OC, device="cuda", dtype=torch.bfloat16)
        dy = torch.randn(B * OH * OH, OC, device="cuda").to(torch.bfloat16)
        dw = torch.zeros(G, ocg, kg, device="cuda")
        line = [name]
        for path, L in libs:
            f = timeit(lambda: L.cxn_conv_fprop(x.data_ptr(), B, H, H, C, cg, G, gran, w.data_ptr(), kg, ocg, k, k, s, pad, pad,
                                                OH, OH, y.data_ptr(), OC, None, 1, st))
            g = timeit(lambda: L.cxn_conv_wgrad(x.data_ptr(), B, H, H, C, cg, G, gran, dy.data_ptr(), OC, ocg, k, k, s, pad, pad,
                                                OH, OH, dw.data_ptr(), kg, 0, st))
            extra = ""
            if hasattr(L, "cxn_conv_tune"):
                L.cxn_conv_tune(0, 1)
                f1 = timeit(lambda: L.cxn_conv_fprop(x.data_ptr(), B, H, H, C, cg, G, gran, w.data_ptr(), kg, ocg, k, k, s, pad,
                                                     pad, OH, OH, y.data_ptr(), OC, None, 1, st))
                L.cxn_conv_tune(0, 2)
                f2 = timeit(lambda: L.cxn_conv_fprop(x.data_ptr(), B, H, H, C, cg, G, gran, w.data_ptr(), kg, ocg, k, k, s, pad,
                                                     pad, OH, OH, y.data_ptr(), OC, None, 1, st))
                L.cxn_conv_tune(0, 0)
                extra = " (mh1 %.1f mh2 %.1f)" % (f1, f2)
            line.append("fprop %.1f wgrad %.1f us%s" % (f, g, extra))
        print(" | ".join(line), flush=True)


if __name__ == "__main__":
    main()
