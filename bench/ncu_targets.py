#!/usr/bin/env python
"""A few launches of the kernels added in round 2, for `ncu --set full` (one GPU): the CTA-pair GEMM on 4096^3, the
CTA-pair convolution on the conv3 forward problem, the eight-epilogue-warp GEMM on the fc6 weight gradient, and the
single-rank forms of the fused exchange and of the sharded-layer collectives."""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch

from cxxnet_b200 import ops

L = ops.lib()
st = ops.stream


def gemm(M, N, K, amn, bmn, f32=False, pair=0):
    A = torch.randn((K, M) if amn else (M, K), device="cuda").to(torch.bfloat16)
    B = torch.randn((K, N) if bmn else (N, K), device="cuda").to(torch.bfloat16)
    D = torch.zeros(M, N, device="cuda", dtype=torch.float32 if f32 else torch.bfloat16)
    L.cxn_gemm_pair(pair)
    for _ in range(3):
        ops.gemm(A.data_ptr(), A.shape[1], 0, amn, B.data_ptr(), B.shape[1], 0, bmn, D.data_ptr(), N, 0, None, 0, M, N, K,
                 flags=0 if f32 else ops.GEMM_OUT_BF16, bn=256 if pair else 0)
    torch.cuda.synchronize()
    L.cxn_gemm_pair(-1)


def conv3(pair):
    N, C, H, OC = 256, 256, 13, 384
    x = torch.randn(N, H, H, C, device="cuda").to(torch.bfloat16)
    w = torch.randn(OC, 9 * C, device="cuda").to(torch.bfloat16) * 0.02
    bias = torch.zeros(OC, device="cuda")
    y = torch.zeros(N * H * H, OC, dtype=torch.bfloat16, device="cuda")
    L.cxn_conv_tune(10, pair)
    for _ in range(3):
        ops.check(L.cxn_conv_fprop(x.data_ptr(), N, H, H, C, C, 1, 8, w.data_ptr(), 9 * C, OC, 3, 3, 1, 1, 1, H, H, y.data_ptr(),
                                   OC, bias.data_ptr(), 1, st()), "fprop")
    torch.cuda.synchronize()
    L.cxn_conv_tune(10, -1)


def exchange_single_rank():
    import test_gpu_exchange_single as T
    lib = T._lib()
    blk = T.Block(lib, [(2048, 2048, 2048)], [dict(lr=0.05, wd=0.01, momentum=0.9)], extra_bytes=1 << 24)
    counter = torch.zeros(4, dtype=torch.int32, device="cuda")
    for call in range(3):
        blk.g.normal_()
        lib.cxn_fused_rs_update_ag(blk.peers, 1, 0, blk.off_w, blk.off_g, blk.off_m, blk.off_m2, blk.off_shadow, blk.off_flags,
                                   blk.segs.data_ptr(), blk.chunk_seg.data_ptr(), 0, blk.total // T.CHUNK, 0, call + 1, 2,
                                   counter.data_ptr(), 296, st())
    torch.cuda.synchronize()
    # sharded-layer collectives with one rank: a 256 x 4096 bf16 block pushed, then pulled
    P, I, LL = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong
    lib.cxn_tp_push.argtypes = [P, I, I, LL, I, P, P, LL, LL, P, I, P]
    lib.cxn_tp_reduce_pull.argtypes = [P, I, I, LL, I, P, LL, LL, LL, I, I, P, LL, I, P]
    buf = torch.zeros(1 << 24, dtype=torch.uint8, device="cuda")
    peers = (ctypes.c_ulonglong * 1)(buf.data_ptr())
    counters = torch.zeros(lib.cxn_tp_counter_bytes() // 4, dtype=torch.int32, device="cuda")
    src = torch.randn(256, 4096, device="cuda").to(torch.bfloat16)
    dst = torch.zeros(256, 4096, dtype=torch.bfloat16, device="cuda")
    desc = (ctypes.c_longlong * 5)(0, 1 << 20, 256, 256, 8192)
    for _ in range(3):
        ops.check(lib.cxn_tp_push(peers, 1, 0, 0, 1, counters.data_ptr(), src.data_ptr(), 8192, 8192, desc, 128, st()), "push")
        ops.check(lib.cxn_tp_reduce_pull(peers, 1, 0, 0, 2, counters.data_ptr(), 1 << 20, 8192, 0, 256, 8192, dst.data_ptr(),
                                         8192, 128, st()), "pull")
    torch.cuda.synchronize()
    assert torch.equal(src, dst)


gemm(4096, 4096, 4096, 0, 0, pair=2)
gemm(4096, 4096, 4096, 0, 0, pair=0)
gemm(4096, 9216, 256, 1, 1, f32=True)          # fc6 weight gradient: eight epilogue warps
conv3(2)
conv3(0)
exchange_single_rank()
print("ok")
