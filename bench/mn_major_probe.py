#!/usr/bin/env python
"""Does the operand layout change the tcgen05 main-loop rate?  4096^3 and a convolution-weight-gradient-like shape
(M = 384, N = 2304, K = 43264, split-K) with K-major and MN-major operands, fp32 and bf16 results."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

from cxxnet_b200 import ops
from gemm_sweep import timeit


def run(M, N, K, amn, bmn, f32, splits=1, bn=0):
    A = torch.randn((K, M) if amn else (M, K), device="cuda").to(torch.bfloat16)
    B = torch.randn((K, N) if bmn else (N, K), device="cuda").to(torch.bfloat16)
    D = torch.zeros(M, N, device="cuda", dtype=torch.float32 if f32 else torch.bfloat16)
    flags = (0 if f32 else ops.GEMM_OUT_BF16) | (ops.GEMM_ATOMIC if splits > 1 else 0)
    t = timeit(lambda: ops.gemm(A.data_ptr(), A.shape[1], 0, amn, B.data_ptr(), B.shape[1], 0, bmn, D.data_ptr(), N, 0, None,
                                0, M, N, K, splits=splits, flags=flags, bn=bn))
    print("M %5d N %5d K %6d  a_mn %d b_mn %d  %s splits %2d bn %3d : %8.1f us  %7.1f TFLOP/s" %
          (M, N, K, amn, bmn, "fp32" if f32 else "bf16", splits, bn, t, 2.0 * M * N * K / t / 1e6))


def main():
    L = ops.lib()
    L.cxn_gemm_pair(0)
    for amn, bmn in [(0, 0), (1, 0), (0, 1), (1, 1)]:
        run(4096, 4096, 4096, amn, bmn, False, bn=256)
    for amn, bmn in [(0, 0), (1, 1)]:
        run(384, 2304, 43264, amn, bmn, True, splits=21, bn=256)
        run(384, 2304, 43264, amn, bmn, True, splits=5, bn=256)
    L.cxn_gemm_pair(-1)


if __name__ == "__main__":
    main()
