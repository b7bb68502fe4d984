#!/usr/bin/env python
"""Micro-benchmark of the tcgen05 GEMM on the AlexNet fully connected shapes (and a square
reference point) against torch.matmul (cuBLAS): time, TFLOP/s and operand bandwidth."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

from cxxnet_b200 import ops


def timeit(fn, iters=20):
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for _ in range(3):
        fn()
    ts = []
    for _ in range(iters):
        flush.max()                         # evict L2 with clean lines (a write would leave dirty lines to drain)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2] * 1e3


def main():
    rows = []
    shapes = [("fc6 fwd b256", 256, 4096, 9216, 0, 0), ("fc6 fwd b32", 32, 4096, 9216, 0, 0),
              ("fc7 fwd b256", 256, 4096, 4096, 0, 0), ("fc8 fwd b256", 256, 1000, 4096, 0, 0),
              ("fc6 dgrad b256", 256, 9216, 4096, 0, 1), ("fc6 wgrad b256", 4096, 9216, 256, 1, 1),
              ("fc6 wgrad b32", 4096, 9216, 32, 1, 1), ("fc6 dgrad b32", 32, 9216, 4096, 0, 1), ("fc7 fwd b32", 32, 4096, 4096, 0, 0), ("square 4096", 4096, 4096, 4096, 0, 0),
              ("conv3-like 43264x384x2304", 43264, 384, 2304, 0, 0),
              ("square 8192", 8192, 8192, 8192, 0, 0)]
    for name, M, N, K, amn, bmn in shapes:
        A = torch.randn((K, M) if amn else (M, K), device="cuda").to(torch.bfloat16)
        B = torch.randn((K, N) if bmn else (N, K), device="cuda").to(torch.bfloat16)
        wgrad = amn and bmn
        D = torch.zeros(M, N, device="cuda", dtype=torch.float32 if wgrad else torch.bfloat16)
        flags = 0 if wgrad else ops.GEMM_OUT_BF16
        res = {"shape": name, "M": M, "N": N, "K": K}
        for bn in (0, 64, 128, 256):
            def run(bn=bn):
                ops.gemm(A.data_ptr(), A.shape[1], 0, amn, B.data_ptr(), B.shape[1], 0, bmn, D.data_ptr(), N, 0, None, 0,
                         M, N, K, flags=flags, bn=bn)
            res[("bn%d_us" % bn) if bn else "auto_us"] = round(timeit(run), 1)
        # CTA pairs (cta_group::2, 256 x 256 tiles), where the shape allows them
        if M > 128 and N > 128 and hasattr(ops.lib(), "cxn_gemm_pair"):
            ops.lib().cxn_gemm_pair(2)
            try:
                def run_pair():
                    ops.gemm(A.data_ptr(), A.shape[1], 0, amn, B.data_ptr(), B.shape[1], 0, bmn, D.data_ptr(), N, 0, None, 0,
                             M, N, K, flags=flags, bn=256)
                res["pair_us"] = round(timeit(run_pair), 1)
            finally:
                ops.lib().cxn_gemm_pair(-1)
        Af = A.t() if amn else A
        Bf = B if bmn else B.t()

        def ref():
            torch.matmul(Af, Bf)
        res["cublas_us"] = round(timeit(ref), 1)
        best = min(res["auto_us"], res["bn64_us"], res["bn128_us"], res["bn256_us"], res.get("pair_us", 1e30))
        res["best_tflops"] = 2.0 * M * N * K / best / 1e6
        res["best_GBps"] = (A.numel() * 2 + B.numel() * 2 + D.numel() * D.element_size()) / best / 1e3
        rows.append(res)
        print(json.dumps(res))
    return rows


if __name__ == "__main__":
    main()
