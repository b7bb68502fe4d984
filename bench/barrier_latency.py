#!/usr/bin/env python
"""Single-rank latency of the barriered kernels (push, pull-reduce, fused exchange) on small payloads: what one
synchronisation point costs before any peer is involved.  CUDA events around 200 back-to-back launches."""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch

from cxxnet_b200 import ops
import test_gpu_exchange_single as T

lib = T._lib()
st = ops.stream
P, I, LL = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong
lib.cxn_tp_push.argtypes = [P, I, I, LL, I, P, P, LL, LL, P, I, P]
lib.cxn_tp_reduce_pull.argtypes = [P, I, I, LL, I, P, LL, LL, LL, I, I, P, LL, I, P]


def timed(fn, n=200):
    for _ in range(10):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3


buf = torch.zeros(1 << 24, dtype=torch.uint8, device="cuda")
peers = (ctypes.c_ulonglong * 1)(buf.data_ptr())
counters = torch.zeros(lib.cxn_tp_counter_bytes() // 4, dtype=torch.int32, device="cuda")
for rows in (32, 256):
    src = torch.randn(rows, 4096, device="cuda").to(torch.bfloat16)
    dst = torch.zeros(rows, 4096, dtype=torch.bfloat16, device="cuda")
    desc = (ctypes.c_longlong * 5)(0, 1 << 20, rows, rows, 8192)
    t_push = timed(lambda: lib.cxn_tp_push(peers, 1, 0, 0, 1, counters.data_ptr(), src.data_ptr(), 8192, 8192, desc, 128, st()))
    t_pull = timed(lambda: lib.cxn_tp_reduce_pull(peers, 1, 0, 0, 2, counters.data_ptr(), 1 << 20, 8192, 0, rows, 8192,
                                                  dst.data_ptr(), 8192, 128, st()))
    print("%d x 4096 bf16 (%d KB): push %.2f us, pull-reduce %.2f us per launch" % (rows, rows * 8, t_push, t_pull))
    assert torch.equal(src, dst)
blk = T.Block(lib, [(64, 1024, 1024)], [dict(lr=0.05, wd=0.01, momentum=0.9)])
counter = torch.zeros(4, dtype=torch.int32, device="cuda")
seq = [0]


def fused():
    seq[0] += 1
    lib.cxn_fused_rs_update_ag(blk.peers, 1, 0, blk.off_w, blk.off_g, blk.off_m, blk.off_m2, blk.off_shadow, blk.off_flags,
                               blk.segs.data_ptr(), blk.chunk_seg.data_ptr(), 0, blk.total // T.CHUNK, 0, seq[0], 2,
                               counter.data_ptr(), 16, st())


print("fused exchange, 64 K parameters: %.2f us per launch" % timed(fused))
