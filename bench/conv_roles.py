#!/usr/bin/env python
"""Where do the warp roles of the implicit-GEMM convolution spend their cycles?  Uses the
kernel's own clock64 counters (cxn_conv_set_prof) - no profiler attached."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

from cxxnet_b200 import ops

L = ops.lib()
B = int(os.environ.get("BATCH", "256"))
SHAPES = [("conv2 fprop", 27, 96, 2, 256, 5, 2), ("conv3 fprop", 13, 256, 1, 384, 3, 1),
          ("conv5 fprop", 13, 384, 2, 256, 3, 1), ("conv2 dgrad", 27, 256, 2, 96, 5, 2)]
NAMES = ["tma wait-empty", "tma total", "mma wait-full", "mma wait-tmem", "mma total", "gather wait-empty",
         "gather setup", "gather total", "epi wait-tmem-full", "epi body", "epi total"]


def main():
    st = torch.cuda.current_stream().cuda_stream
    prof = torch.zeros(148, 16, dtype=torch.int64, device="cuda")
    modes = [("normal", 0), ("all-lane wait", 64), ("no gather/tma/mma", 7), ("no gather/tma/mma all-lane", 7 + 64)]
    for name, H, C, G, OC, k, pad in SHAPES:
        cg, ocg = C // G, OC // G
        x = torch.randn(B, H, H, C, device="cuda").to(torch.bfloat16)
        w = (torch.randn(G, ocg, k, k, cg, device="cuda") * 0.05).to(torch.bfloat16)
        y = torch.empty(B * H * H, OC, device="cuda", dtype=torch.bfloat16)
        for mname, dbg in modes:
            L.cxn_conv_tune(0, 1)
            L.cxn_conv_tune(5, dbg)

            def run():
                return L.cxn_conv_fprop(x.data_ptr(), B, H, H, C, cg, G, 8, w.data_ptr(), k * k * cg, ocg, k, k, 1, pad, pad,
                                        H, H, y.data_ptr(), OC, None, 0, st)
            for _ in range(3):
                run()
            torch.cuda.synchronize()
            prof.zero_()
            L.cxn_conv_set_prof(prof.data_ptr())
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            run()
            e1.record()
            torch.cuda.synchronize()
            L.cxn_conv_set_prof(None)
            us = e0.elapsed_time(e1) * 1e3
            m = prof.double().mean(dim=0).cpu().tolist()
            print("%s [%s] %.1f us; mean cycles per CTA: %s" % (name, mname, us, ", ".join("%s %.0fk" % (n, v / 1e3) for n, v in zip(NAMES, m))))
    L.cxn_conv_tune(0, 0)
    L.cxn_conv_tune(5, 0)


if __name__ == "__main__":
    main()
