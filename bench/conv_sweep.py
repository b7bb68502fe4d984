#!/usr/bin/env python
"""A/B sweep of the implicit-GEMM convolution's tuning knobs (cxn_conv_tune) on the AlexNet
convolution shapes: forward, the stride-1 backward-data problems (same kernel on dy) and the
weight gradient, each checked against PyTorch on the same bf16 inputs."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.nn.functional as F

from cxxnet_b200 import ops

L = ops.lib()
B = int(os.environ.get("BATCH", "256"))
# name, H, C, G, OC, k, pad
SHAPES = [("conv2 fprop", 27, 96, 2, 256, 5, 2), ("conv3 fprop", 13, 256, 1, 384, 3, 1),
          ("conv4 fprop", 13, 384, 2, 384, 3, 1), ("conv5 fprop", 13, 384, 2, 256, 3, 1),
          ("conv2 dgrad", 27, 256, 2, 96, 5, 2), ("conv3 dgrad", 13, 384, 1, 256, 3, 1),
          ("conv4 dgrad", 13, 384, 2, 384, 3, 1), ("conv5 dgrad", 13, 256, 2, 384, 3, 1)]
CONFIGS = [("gather 128-row", dict(atma=0, mh=1, stages=-1)),
           ("gather 256-row", dict(atma=0, mh=2, bn=128)),
           ("tma 128-row", dict(atma=1, mh=1, stages=-1)),
           ("tma 256x128", dict(atma=1, mh=2, stages=4, bn=128)),
           ("tma 256x256", dict(atma=1, mh=2, stages=3, bn=256)),
           ("auto", dict(atma=-1, mh=0, stages=0)),
           # opt-in routes (EXTRA=1): CTA pairs sharing the filter tile by multicast; narrower tiles for small grids
           ("auto + cluster", dict(atma=-1, mh=0, stages=0, cluster=2, extra=1)),
           ("tma 128-row + cluster", dict(atma=1, mh=1, stages=-1, cluster=2, extra=1)),
           ("auto + fill", dict(atma=-1, mh=0, stages=0, fill=1, extra=1)),
           ("cta pairs", dict(atma=-1, mh=0, stages=0, pair=2, extra=1)),
           ("halo 2 halves", dict(atma=-1, mh=0, stages=0, halo=1, extra=1)),
           ("halo 1 half", dict(atma=-1, mh=0, stages=0, halo=2, extra=1))]
if os.environ.get("EXTRA", "0") != "1":
    CONFIGS = [c for c in CONFIGS if not c[1].get("extra")]
WCONFIGS = [("gather", dict(atma=0)), ("tma", dict(atma=1))]


def set_knobs(c, ocg):
    L.cxn_conv_tune(0, c.get("mh", 0))
    st = c.get("stages", 0)
    if st == -1:
        st = 4 if ocg > 128 else 6
    L.cxn_conv_tune(1, st)
    L.cxn_conv_tune(4, c.get("bn", 0))
    L.cxn_conv_tune(5, c.get("dbg", 0))
    L.cxn_conv_tune(6, c.get("atma", -1))
    L.cxn_conv_tune(7, c.get("cluster", 0))
    L.cxn_conv_tune(8, c.get("fill", 0))
    L.cxn_conv_tune(9, c.get("halo", 0))
    L.cxn_conv_tune(10, c.get("pair", 0))


def reset_knobs():
    for k_ in range(7):
        L.cxn_conv_tune(k_, 0 if k_ in (0, 1, 4, 5) else -1)
    L.cxn_conv_tune(7, 0)
    L.cxn_conv_tune(8, 0)
    L.cxn_conv_tune(9, 0)
    L.cxn_conv_tune(10, 0)


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
    torch.manual_seed(0)
    st = torch.cuda.current_stream().cuda_stream
    which = os.environ.get("WHICH", "fprop,wgrad")
    for name, H, C, G, OC, k, pad in SHAPES:
        cg, ocg = C // G, OC // G
        x = torch.randn(B, H, H, C, device="cuda").to(torch.bfloat16)
        w = (torch.randn(G, ocg, k, k, cg, device="cuda") * 0.05).to(torch.bfloat16)
        y = torch.empty(B * H * H, OC, device="cuda", dtype=torch.bfloat16)
        wt = w.reshape(OC, k, k, cg).permute(0, 3, 1, 2).float()
        flops = 2.0 * B * H * H * OC * k * k * cg
        if "fprop" in which:
            ref = F.conv2d(x.permute(0, 3, 1, 2).float(), wt, padding=pad, groups=G).permute(0, 2, 3, 1).reshape(B * H * H, OC)
            out = {"shape": name}
            for cname, c in CONFIGS:
                set_knobs(c, ocg)

                def run():
                    return L.cxn_conv_fprop(x.data_ptr(), B, H, H, C, cg, G, 8, w.data_ptr(), k * k * cg, ocg, k, k, 1, pad,
                                            pad, H, H, y.data_ptr(), OC, None, 0, st)
                y.zero_()
                rc = run()
                if rc != 0:
                    out[cname] = "rc=%d" % rc
                    continue
                torch.cuda.synchronize()
                err = float((y.float() - ref).abs().max() / ref.abs().max())
                us = timeit(run)
                out[cname] = "%.1f us %.0f TF err %.1e" % (us, flops / us / 1e6, err)
            print(json.dumps(out, indent=0), flush=True)
        if "wgrad" in which and "fprop" in name:
            # weight gradient of the same layer: dy random
            dy = torch.randn(B * H * H, OC, device="cuda").to(torch.bfloat16)
            kg = k * k * cg
            dw = torch.zeros(G, ocg, kg, device="cuda", dtype=torch.float32)
            xr = x.permute(0, 3, 1, 2).float().requires_grad_(False)
            wr = wt.clone().requires_grad_(True)
            yr = F.conv2d(xr, wr, padding=pad, groups=G)
            yr.backward(dy.float().reshape(B, H, H, OC).permute(0, 3, 1, 2))
            refw = wr.grad.reshape(G, ocg, cg, k, k).permute(0, 1, 3, 4, 2).reshape(G, ocg, kg)
            out = {"shape": name.replace("fprop", "wgrad")}
            for cname, c in WCONFIGS:
                set_knobs(c, ocg)

                def runw():
                    return L.cxn_conv_wgrad(x.data_ptr(), B, H, H, C, cg, G, 8, dy.data_ptr(), OC, ocg, k, k, 1, pad, pad, H, H,
                                            dw.data_ptr(), kg, 0, st)
                dw.zero_()
                rc = runw()
                if rc != 0:
                    out[cname] = "rc=%d" % rc
                    continue
                torch.cuda.synchronize()
                err = float((dw - refw).abs().max() / refw.abs().max())
                us = timeit(runw)
                out[cname] = "%.1f us %.0f TF err %.1e" % (us, flops / us / 1e6, err)
            print(json.dumps(out, indent=0), flush=True)
    reset_knobs()


if __name__ == "__main__":
    main()
