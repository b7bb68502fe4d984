#!/usr/bin/env python
"""Bandwidth-bound kernels of the AlexNet step, one by one, against their byte roofline
(algorithmic bytes / measured copy bandwidth from MEASURED_PEAKS.json, fallback 6.5 TB/s)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

from cxxnet_b200 import ops

L = ops.lib()
B = int(os.environ.get("BATCH", "256"))


def peak_gbps():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            d = json.load(f)
        for k in ("hbm_gbs", "hbm_copy_GBps", "copy_GBps"):
            if k in d:
                return float(d[k])
    except Exception:
        pass
    return 6500.0


def timeit(run, iters=10):
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for _ in range(2):
        run()
    ts = []
    for _ in range(iters):
        flush.max()                         # evict L2 with clean lines (a write would leave 126 MB of dirty lines to drain)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        run()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def report(name, us, nbytes, peak):
    print("%-34s %8.1f us  %7.1f MB  %6.0f GB/s  %4.0f%% of copy peak" % (name, us, nbytes / 1e6, nbytes / us / 1e3,
                                                                       100.0 * nbytes / us / 1e3 / peak), flush=True)


def main():
    st = torch.cuda.current_stream().cuda_stream
    peak = peak_gbps()
    bf = torch.bfloat16
    # pooling: (name, H, C, k, stride, relu)
    for name, H, C, k, s, relu in [("pool1", 55, 96, 3, 2, 1), ("pool2", 27, 256, 3, 2, 1), ("pool5", 13, 256, 3, 2, 1)]:
        OH = min(H - k + s - 1, H - 1) // s + 1
        x = torch.randn(B, H, H, C, device="cuda").to(bf)
        y = torch.empty(B, OH, OH, C, device="cuda", dtype=bf)
        dy = torch.randn(B, OH, OH, C, device="cuda").to(bf)
        dx = torch.empty_like(x)
        us = timeit(lambda: L.cxn_pool_fwd(x.data_ptr(), y.data_ptr(), B, H, H, C, k, k, s, OH, OH, 0, relu, st))
        report(name + " fwd", us, (x.numel() + y.numel()) * 2, peak)
        us = timeit(lambda: L.cxn_pool_bwd(x.data_ptr(), y.data_ptr(), dy.data_ptr(), dx.data_ptr(), B, H, H, C, k, k, s, OH, OH,
                                           0, relu, st))
        report(name + " bwd", us, (2 * x.numel() + 2 * y.numel()) * 2, peak)
    # LRN
    for name, H, C in [("norm1", 27, 96), ("norm2", 13, 256)]:
        x = torch.randn(B * H * H, C, device="cuda").to(bf)
        y = torch.empty_like(x)
        dy = torch.randn_like(x)
        dx = torch.empty_like(x)
        us = timeit(lambda: L.cxn_lrn_fwd(x.data_ptr(), y.data_ptr(), B * H * H, C, 5, 1e-4, 0.75, 1.0, st))
        report(name + " lrn fwd", us, 2 * x.numel() * 2, peak)
        us = timeit(lambda: L.cxn_lrn_bwd(x.data_ptr(), dy.data_ptr(), dx.data_ptr(), B * H * H, C, 5, 1e-4, 0.75, 1.0, st))
        report(name + " lrn bwd", us, 3 * x.numel() * 2, peak)
    # bias gradient = column sums of dy
    for name, rows, C in [("conv1 bias grad", B * 55 * 55, 96), ("conv2 bias grad", B * 27 * 27, 256), ("conv3 bias grad", B * 169, 384)]:
        dy = torch.randn(rows, C, device="cuda").to(bf)
        out = torch.zeros(C, device="cuda")
        us = timeit(lambda: L.cxn_colsum_bf16(dy.data_ptr(), rows, C, C, out.data_ptr(), 0, st))
        report(name, us, dy.numel() * 2, peak)
    # input layout conversion
    x = torch.randn(B, 3, 227, 227, device="cuda")
    y = torch.empty(B, 227, 227, 4, device="cuda", dtype=bf)
    us = timeit(lambda: L.cxn_nchw_f32_to_nhwc_bf16(x.data_ptr(), y.data_ptr(), B, 3, 227 * 227, 4, st))
    report("input nchw f32 -> nhwc4 bf16", us, x.numel() * 4 + y.numel() * 2, peak)


if __name__ == "__main__":
    main()
