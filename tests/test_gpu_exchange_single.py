"""The peer-memory exchange kernels (csrc/kernels/fused_exchange.cu) driven as a ONE-rank exchange, so that a
single-GPU `pytest -m gpu` run executes them: `fused_rs_update_ag_kernel` against the closed-form sgd / nag / adam
step in plain fp32 PyTorch (reference formulas: src/updater/sgd_updater-inl.hpp:73-84, nag_updater-inl.hpp,
adam_updater-inl.hpp), buckets with more segments than the kernel caches in shared memory, and the all-gather /
broadcast kernels.  The multi-rank behaviour of the same kernels is tests/test_gpu_multi.py (needs 2 GPUs)."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

SEG = np.dtype([("start", "<i8"), ("n", "<i8"), ("shadow_off", "<i8"), ("cols", "<i4"), ("shadow_ld", "<i4"),
                ("lr", "<f4"), ("wd", "<f4"), ("momentum", "<f4"), ("clip", "<f4"),
                ("adam_lr_t", "<f4"), ("beta1", "<f4"), ("beta2", "<f4"), ("pad", "<f4")])
CHUNK = 1024
KIND = {"sgd": 0, "nag": 1, "adam": 2}


def _lib():
    from cxxnet_b200 import ops
    L = ops.lib()
    P, I, LL, U = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_uint
    L.cxn_fused_rs_update_ag.argtypes = [P, I, I, LL, LL, LL, LL, LL, LL, P, P, LL, LL, I, U, I, P, I, P]
    L.cxn_fused_rs_update_ag.restype = I
    L.cxn_symm_broadcast.argtypes = [P, I, I, LL, LL, LL, I, U, I, P, I, P]
    L.cxn_symm_broadcast.restype = I
    L.cxn_symm_allgather.argtypes = [P, I, I, LL, P, LL, LL, LL, P, LL, LL, LL, I, U, I, P, I, P]
    L.cxn_symm_allgather.restype = I
    L.cxn_xchg_set_epoch_ptr.argtypes = [P]
    L.cxn_xchg_set_epoch_ptr(None)
    return L


class Block:
    """one rank's block laid out by hand: flags | w | g | m | m2 | shadow | scratch"""

    def __init__(self, lib, shapes, hyper, extra_bytes=0):
        flag_bytes = lib.cxn_xchg_flag_bytes()
        starts, total, soffs, stotal = [], 0, [], 0
        for rows, cols, sld in shapes:
            starts.append(total)
            total += (rows * cols + CHUNK - 1) // CHUNK * CHUNK
            soffs.append(stotal if sld else -1)
            if sld:
                stotal += (rows * sld + 127) // 128 * 128
        self.total, self.starts, self.shapes, self.soffs = total, starts, shapes, soffs
        r256 = lambda x: (x + 255) // 256 * 256
        self.off_flags = 0
        self.off_w = r256(flag_bytes)
        self.off_g = self.off_w + total * 4
        self.off_m = self.off_g + total * 4
        self.off_m2 = self.off_m + total * 4
        self.off_shadow = self.off_m2 + total * 4
        self.off_extra = r256(self.off_shadow + stotal * 2 + 256)
        self.buf = torch.zeros(self.off_extra + extra_bytes, dtype=torch.uint8, device="cuda")
        f32 = lambda off: self.buf[off:off + total * 4].view(torch.float32)
        self.w, self.g, self.m, self.m2 = f32(self.off_w), f32(self.off_g), f32(self.off_m), f32(self.off_m2)
        self.shadow = self.buf[self.off_shadow:self.off_shadow + stotal * 2].view(torch.bfloat16)
        segs = np.zeros(len(shapes), dtype=SEG)
        cs = np.zeros(total // CHUNK, dtype=np.int32)
        for i, ((rows, cols, sld), st) in enumerate(zip(shapes, starts)):
            h = hyper[i % len(hyper)]
            n = rows * cols
            segs[i] = (st, n, soffs[i], cols, sld, h["lr"], h["wd"], h["momentum"], h.get("clip", 0.0),
                       h.get("adam_lr_t", 0.0), h.get("beta1", 0.1), h.get("beta2", 0.001), 0.0)
            cs[st // CHUNK:(st + (n + CHUNK - 1) // CHUNK * CHUNK) // CHUNK] = i
        self.hyper = [hyper[i % len(hyper)] for i in range(len(shapes))]
        self.segs = torch.from_numpy(np.frombuffer(segs.tobytes(), dtype=np.uint8).copy()).cuda()
        self.chunk_seg = torch.from_numpy(cs).cuda()
        self.peers = (ctypes.c_ulonglong * 1)(self.buf.data_ptr())

    def seg(self, t, i):
        rows, cols, _ = self.shapes[i]
        return t[self.starts[i]:self.starts[i] + rows * cols]


def _closed_form(kind, w, g, m, m2, h):
    """one update in fp32 torch, the arithmetic of the reference updaters"""
    lr, wd, mom = h["lr"], h["wd"], h["momentum"]
    if kind == "sgd":
        if h.get("clip", 0.0) != 0.0:
            g = torch.nan_to_num(g, nan=0.0).clamp(-h["clip"], h["clip"])
        m = mom * m - lr * (g + wd * w)
        w = w + m
    elif kind == "nag":
        old = m
        m = mom * old - lr * (g + wd * w)
        w = w + (1.0 + mom) * m - mom * old
    else:
        if wd > 0.0:
            g = g - wd * w
        m = m + h["beta1"] * (g - m)
        m2 = m2 + h["beta2"] * (g * g - m2)
        w = w - h["adam_lr_t"] * (m / (m2.sqrt() + 1e-8))
    return w, m, m2


@pytest.mark.parametrize("kind", ["sgd", "nag", "adam"])
@pytest.mark.parametrize("nsegs", [3, 21])
def test_fused_exchange_single_rank_matches_closed_form(kind, nsegs):
    """world = 1: the reduce-scatter degenerates to a copy, everything else (updater arithmetic, gradient zeroing,
    bf16 shadow refresh, both flag barriers) runs as on several GPUs.  21 segments > the 16 the kernel caches in
    shared memory: the tail segments come from the global table and must be updated like the rest."""
    lib = _lib()
    base = [(37, 50, 56), (1, 300, 0), (4, 1024, 1024), (1, 5, 0), (3, 7, 8)]
    shapes = [base[i % len(base)] for i in range(nsegs)]
    hyper = [dict(lr=0.05, wd=0.01, momentum=0.9, clip=0.5, adam_lr_t=0.02, beta1=0.1, beta2=0.001),
             dict(lr=0.02, wd=0.0, momentum=0.5, adam_lr_t=0.01, beta1=0.2, beta2=0.01)]
    blk = Block(lib, shapes, hyper)
    torch.manual_seed(5 + nsegs)
    blk.w.copy_(torch.randn(blk.total, device="cuda"))
    ref = [(blk.seg(blk.w, i).clone(), torch.zeros_like(blk.seg(blk.w, i)), torch.zeros_like(blk.seg(blk.w, i)))
           for i in range(nsegs)]
    counter = torch.zeros(4, dtype=torch.int32, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    for call in range(3):
        g = torch.randn(blk.total, device="cuda")
        if kind == "sgd" and call == 1:
            g[blk.starts[0] + 3] = float("nan")          # the sgd updater clips, and maps NaN to 0 (segment 0 clips)
        blk.g.copy_(g)
        rc = lib.cxn_fused_rs_update_ag(blk.peers, 1, 0, blk.off_w, blk.off_g, blk.off_m, blk.off_m2, blk.off_shadow,
                                        blk.off_flags, blk.segs.data_ptr(), blk.chunk_seg.data_ptr(), 0,
                                        blk.total // CHUNK, KIND[kind], call + 1, 2, counter.data_ptr(), 8, st)
        assert rc == 0
        torch.cuda.synchronize()
        for i in range(nsegs):
            w, m, m2 = _closed_form(kind, ref[i][0], blk.seg(g, i), ref[i][1], ref[i][2], blk.hyper[i])
            ref[i] = (w, m, m2)
            assert torch.allclose(blk.seg(blk.w, i), w, atol=1e-5, rtol=1e-5), (kind, call, i)
            assert float(blk.seg(blk.g, i).abs().max()) == 0.0, "gradient not consumed (segment %d)" % i
            rows, cols, sld = shapes[i]
            if sld:
                sh = blk.shadow[blk.soffs[i]:blk.soffs[i] + rows * sld].view(rows, sld)[:, :cols]
                assert torch.equal(sh, blk.seg(blk.w, i).view(rows, cols).to(torch.bfloat16)), (kind, call, i)
    assert int(counter[0]) == 0


def test_allgather_and_broadcast_single_rank():
    """world = 1: the all-gather places this rank's pieces (zero-filling past the valid bytes), the broadcast from
    the only rank leaves the block untouched; both pass their barriers"""
    lib = _lib()
    blk = Block(lib, [(1, 1024, 0)], [dict(lr=0.1, wd=0.0, momentum=0.0)], extra_bytes=1 << 16)
    counter = torch.zeros(4, dtype=torch.int32, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    src0 = torch.randn(4096, device="cuda").to(torch.bfloat16)
    src1 = torch.randn(1024, device="cuda").to(torch.bfloat16)
    dst = blk.buf[blk.off_extra:]
    dst.fill_(0xAB)
    o0, o1 = blk.off_extra, blk.off_extra + 16384
    rc = lib.cxn_symm_allgather(blk.peers, 1, 0, blk.off_flags, src0.data_ptr(), o0, 8192, 8192 - 1024,
                                src1.data_ptr(), o1, 2048, 2048, 2, 1, 5, counter.data_ptr(), 8, st)
    assert rc == 0
    torch.cuda.synchronize()
    got0 = dst[:8192].view(torch.bfloat16)
    assert torch.equal(got0[:3584], src0[:3584]) and float(got0[3584:].float().abs().max()) == 0.0
    assert torch.equal(dst[16384:16384 + 2048].view(torch.bfloat16), src1)
    assert int(dst[8192]) == 0xAB                         # nothing written outside the slots
    before = blk.buf.clone()
    rc = lib.cxn_symm_broadcast(blk.peers, 1, 0, blk.off_flags, blk.off_w, 4096, 0, 2, 6, counter.data_ptr(), 8, st)
    assert rc == 0
    torch.cuda.synchronize()
    assert torch.equal(before[blk.off_w:], blk.buf[blk.off_w:])
    assert int(counter[0]) == 0
