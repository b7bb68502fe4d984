"""`cxn_fused_rs_update_ag_noalloc`: the fused gradient-exchange entry point for hosts that must
not allocate device memory.

Who needs it: a process that gives the whole GPU to its own memory pool when it starts (any
`cudaMalloc` issued later by a library fails), or a latency-critical serving / training loop that
cannot tolerate the device-wide synchronisation `cudaMalloc` and `cudaFree` imply.  The regular
entry point takes its one word of per-bucket scratch (the CTA completion counter) from the
caller - i.e. the caller had to allocate it; this variant keeps 256 such words (1 KiB) inside
the library image instead and refuses bucket slots beyond that.

The test drives both entry points through the plain C ABI (ctypes on the shared library, raw
device pointers), as a single-rank exchange so one GPU is enough, and requires identical
results on two consecutive calls - identical to each other and to the sgd step written out in plain
fp32 PyTorch.  It lives in its own file, collected last, so that a
`pytest -x` run has already executed every other test if this one fails; it does not depend on
state left by them (fresh buffers, and the entry point clears its scratch word before use).
"""
import ctypes
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "cxxnet_b200", "lib", "libcxxnet_b200.so")

SEG = np.dtype([("start", "<i8"), ("n", "<i8"), ("shadow_off", "<i8"), ("cols", "<i4"), ("shadow_ld", "<i4"),
                ("lr", "<f4"), ("wd", "<f4"), ("momentum", "<f4"), ("clip", "<f4"),
                ("adam_lr_t", "<f4"), ("beta1", "<f4"), ("beta2", "<f4"), ("pad", "<f4")])
CHUNK = 1024


class Block:
    """one rank's parameter block laid out by hand: flags | w | g | m | shadow"""

    def __init__(self, lib, sizes):
        flag_bytes = lib.cxn_xchg_flag_bytes()
        self.nseg = len(sizes)
        starts, total = [], 0
        for n in sizes:
            starts.append(total)
            total += (n + CHUNK - 1) // CHUNK * CHUNK
        self.total, self.starts, self.sizes = total, starts, sizes
        self.off_flags = 0
        self.off_w = (flag_bytes + 255) // 256 * 256
        self.off_g = self.off_w + total * 4
        self.off_m = self.off_g + total * 4
        self.off_shadow = self.off_m + total * 4
        shadow_elems = sum((n + 127) // 128 * 128 for n in sizes)
        self.buf = torch.zeros(self.off_shadow + shadow_elems * 2 + 256, dtype=torch.uint8, device="cuda")
        f32 = lambda off: self.buf[off:off + total * 4].view(torch.float32)
        self.w, self.g, self.m = f32(self.off_w), f32(self.off_g), f32(self.off_m)
        self.shadow = self.buf[self.off_shadow:self.off_shadow + shadow_elems * 2].view(torch.bfloat16)
        segs = np.zeros(len(sizes), dtype=SEG)
        cs = np.zeros(total // CHUNK, dtype=np.int32)
        soff = 0
        for i, (st, n) in enumerate(zip(starts, sizes)):
            segs[i] = (st, n, soff, n, n, 0.05, 0.001, 0.9, 0.0, 0.0, 0.1, 0.001, 0.0)
            cs[st // CHUNK:(st + (n + CHUNK - 1) // CHUNK * CHUNK) // CHUNK] = i
            soff += (n + 127) // 128 * 128
        self.segs = torch.from_numpy(np.frombuffer(segs.tobytes(), dtype=np.uint8).copy()).cuda()
        self.chunk_seg = torch.from_numpy(cs).cuda()
        self.peers = (ctypes.c_ulonglong * 1)(self.buf.data_ptr())

    def args(self, seq, slot):
        return [self.peers, 1, 0, self.off_w, self.off_g, self.off_m, 0, self.off_shadow, self.off_flags,
                self.segs.data_ptr(), self.chunk_seg.data_ptr(), 0, self.total // CHUNK, 0, seq, slot]


def test_no_alloc_entry_point_matches_regular_entry_point():
    assert torch.cuda.is_available()
    lib = ctypes.CDLL(LIB)
    P, I, LL, U = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_uint
    common = [P, I, I, LL, LL, LL, LL, LL, LL, P, P, LL, LL, I, U, I]
    lib.cxn_fused_rs_update_ag.argtypes = common + [P, I, P]
    lib.cxn_fused_rs_update_ag.restype = I
    lib.cxn_fused_rs_update_ag_noalloc.argtypes = common + [I, P]
    lib.cxn_fused_rs_update_ag_noalloc.restype = I
    lib.cxn_fused_rs_update_ag_noalloc_max_slots.restype = I
    max_slots = lib.cxn_fused_rs_update_ag_noalloc_max_slots()
    assert max_slots == 256

    sizes = [5000, 37, 3 * 1024]
    a, b = Block(lib, sizes), Block(lib, sizes)
    torch.manual_seed(11)
    w0 = torch.randn(a.total, device="cuda")
    a.w.copy_(w0)
    b.w.copy_(w0)
    counter = torch.zeros(4, dtype=torch.int32, device="cuda")       # the scratch the regular entry point needs
    stream = torch.cuda.current_stream().cuda_stream
    slot = 3
    w_ref, m_ref = w0.clone(), torch.zeros_like(w0)                  # closed-form sgd: lr 0.05, wd 0.001, momentum 0.9
    for call in range(2):                                            # two calls in a row
        g = torch.randn(a.total, device="cuda")
        m_ref = 0.9 * m_ref - 0.05 * (g + 0.001 * w_ref)
        w_ref = w_ref + m_ref
        a.g.copy_(g)
        b.g.copy_(g)
        seq = call + 1
        rc = lib.cxn_fused_rs_update_ag(*a.args(seq, slot), counter.data_ptr(), 16, stream)
        assert rc == 0
        rc = lib.cxn_fused_rs_update_ag_noalloc(*b.args(seq, slot), 16, stream)
        assert rc == 0
        torch.cuda.synchronize()
        for st, n in zip(a.starts, sizes):
            assert torch.allclose(a.w[st:st + n], b.w[st:st + n], atol=1e-6, rtol=0), "weights differ (call %d)" % call
            assert torch.allclose(a.m[st:st + n], b.m[st:st + n], atol=1e-6, rtol=0), "momentum differs (call %d)" % call
            assert float(a.g[st:st + n].abs().max()) == 0.0 and float(b.g[st:st + n].abs().max()) == 0.0  # consumed
            # both entry points against the update written out in plain fp32 PyTorch
            assert torch.allclose(b.w[st:st + n], w_ref[st:st + n], atol=1e-5, rtol=1e-5), "closed form (call %d)" % call
            assert torch.allclose(a.w[st:st + n], w_ref[st:st + n], atol=1e-5, rtol=1e-5)
        assert torch.equal(a.shadow, b.shadow)
        assert not torch.equal(b.w, w0)                                             # and something happened
    assert int(counter[0]) == 0                                      # the regular entry point left its scratch zeroed
    # bucket slots beyond the built-in scratch are refused, nothing is launched
    before = b.w.clone()
    rc = lib.cxn_fused_rs_update_ag_noalloc(*b.args(3, max_slots), 16, stream)
    assert rc != 0
    torch.cuda.synchronize()
    assert torch.equal(before, b.w)
