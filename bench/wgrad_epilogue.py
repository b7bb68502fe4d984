#!/usr/bin/env python
"""Fully connected weight-gradient GEMMs (K = batch rows, fp32 result): the 32-column-chunk epilogue against the
WIDE one (whole rows of a warp's columns per store instruction), each with and without the fused updater
(cxn_gemm_wgrad_update), next to torch.matmul (cuBLAS) and to GEMM + multi-tensor update.

  CXN_GEMM_WIDE=0 python bench/wgrad_epilogue.py ; CXN_GEMM_WIDE=1 python bench/wgrad_epilogue.py
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

from cxxnet_b200 import ops
from gemm_sweep import timeit


def main():
    L = ops.lib()
    wide = os.environ.get("CXN_GEMM_WIDE", "1")
    print("# CXN_GEMM_WIDE=%s   us: gemm (fp32 out) | cuBLAS bf16 matmul | gemm + fused updater | bytes/param 4 vs 18" % wide)
    hyper = torch.tensor([0.01, 0.0005, 0.9, 0.0, 0.0, 0.0, 0.0], dtype=torch.float32, device="cuda")
    for name, M, N, K in [("fc6 b256", 4096, 9216, 256), ("fc7 b256", 4096, 4096, 256), ("fc8 b256", 1000, 4096, 256),
                          ("fc6 b32", 4096, 9216, 32), ("fc6 rows/8 b256 (8-GPU shard)", 512, 9216, 256)]:
        ldy = ops.round_up(M, 8)
        dy = torch.zeros(K, ldy, device="cuda", dtype=torch.bfloat16)
        dy[:, :M] = torch.randn(K, M, device="cuda").to(torch.bfloat16)
        x = torch.randn(K, N, device="cuda").to(torch.bfloat16)
        g = torch.zeros(M, N, device="cuda")
        w = torch.randn(M, N, device="cuda")
        m = torch.zeros(M, N, device="cuda")
        sh = torch.zeros(M, N, device="cuda", dtype=torch.bfloat16)
        t_g = timeit(lambda: ops.gemm(dy.data_ptr(), ldy, 0, 1, x.data_ptr(), N, 0, 1, g.data_ptr(), N, 0, None, 0, M, N, K))
        dyt = dy[:, :M].t().contiguous()
        t_c = timeit(lambda: torch.matmul(dyt, x))
        t_f = timeit(lambda: ops.check(L.cxn_gemm_wgrad_update(dy.data_ptr(), ldy, x.data_ptr(), N, M, N, K, w.data_ptr(),
                                                               m.data_ptr(), None, sh.data_ptr(), N, N, hyper.data_ptr(), 0,
                                                               ops.stream()), "fused"))
        n = M * N
        print("%-32s %7.1f (%.2f TB/s) | %7.1f | %7.1f (%.2f TB/s)" % (name, t_g, n * 4 / t_g / 1e6, t_c, t_f, n * 18 / t_f / 1e6))


if __name__ == "__main__":
    main()
