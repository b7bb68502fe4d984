"""Host side of the parameter arenas (cxxnet_b200/parallel/arena.py) without a GPU: segment layout on update-chunk
boundaries, the chunk -> segment map the kernels index, and the per-step hyper-parameter table (schedules, per-tag
values, adam's bias-corrected rate, the no-zero flag) that is staged for the multi-tensor update / exchange kernels."""
import os

import numpy as np
import pytest
import torch

from cxxnet_b200 import ops
from cxxnet_b200.updaters import UpdaterParam

pytestmark = pytest.mark.skipif(not ops.available(), reason="kernel library not built")


def _param(tag, **kv):
    p = UpdaterParam(tag)
    for k, v in kv.items():
        p.set_param(k.replace("__", ":"), str(v))
    return p


def _arena(kind="sgd"):
    from cxxnet_b200.parallel.arena import ParamArena
    a = ParamArena(torch.device("cpu"), kind)
    s0 = a.add(None, 0, "wmat", 96, 363, 368, _param("wmat", lr=0.01, wd=0.0005, lr__schedule="expdecay",
                                                       lr__gamma=0.1, lr__step=100))
    s1 = a.add(None, 0, "bias", 1, 96, 0, _param("bias", lr=0.02, wd=0.0))
    s2 = a.add(None, 3, "wmat", 10, 4096, 4096, _param("wmat", lr=0.01, momentum=0.5, clip_gradient=2.0))
    s1.nozero = True
    a.upload_hyper = lambda slot=0: None          # the device half (pinned table -> device) needs a GPU
    a.finalize()
    return a, (s0, s1, s2)


def test_segments_sit_on_update_chunks_and_map_back():
    from cxxnet_b200.parallel.arena import SEG_DTYPE
    a, (s0, s1, s2) = _arena()
    ch = a.chunk
    assert ch > 0 and all(s.start % ch == 0 for s in (s0, s1, s2))
    assert s0.start == 0 and s1.start == ops.round_up(96 * 363, ch) and s2.start == s1.start + ops.round_up(96, ch)
    assert a.total == s2.start + ops.round_up(10 * 4096, ch) and a.nchunks == a.total // ch
    cs = a.chunk_seg.numpy()
    assert cs.shape[0] == a.nchunks
    for i, s in enumerate((s0, s1, s2)):
        lo, hi = s.start // ch, (s.start + ops.round_up(s.n, ch)) // ch
        assert (cs[lo:hi] == i).all()
    assert a.chunk_range(0, 2) == (0, s2.start // ch) and a.chunk_range(2, 3) == (s2.start // ch, a.nchunks)
    # shadow (bf16, GEMM-ready) rows are padded to their leading dimension; vectors have none
    assert s0.shadow_off == 0 and s1.shadow_off == -1 and s2.shadow_off == ops.round_up(96 * 368, 128)
    assert a.w.numel() >= a.total and a.shadow.numel() >= s2.shadow_off + 10 * 4096
    assert a.w_view(s1).numel() == 96 and a.shadow_view(s1) is None
    assert SEG_DTYPE.itemsize == ops.lib().cxn_segment_bytes()


def test_hyper_table_follows_schedules_and_tags():
    from cxxnet_b200.parallel.arena import SEG_DTYPE
    a, (s0, s1, s2) = _arena()
    assert a.prepare_hyper(0, 0) is True
    t = np.frombuffer(a.seg_host[0].numpy().tobytes(), dtype=SEG_DTYPE)
    assert t[0]["start"] == 0 and t[0]["n"] == 96 * 363 and t[0]["cols"] == 363 and t[0]["shadow_ld"] == 368
    assert np.isclose(t[0]["lr"], 0.01) and np.isclose(t[0]["wd"], 0.0005) and np.isclose(t[0]["momentum"], 0.9)
    assert np.isclose(t[1]["lr"], 0.02) and t[1]["wd"] == 0.0 and t[1]["flags"] == 1 and t[0]["flags"] == 0
    assert np.isclose(t[2]["momentum"], 0.5) and np.isclose(t[2]["clip"], 2.0) and t[2]["shadow_off"] == s2.shadow_off
    assert a.prepare_hyper(0, 0) is False                     # unchanged table: the pinned buffer is left alone
    assert a.prepare_hyper(250, 1) is True                    # expdecay: 0.01 * 0.1 ** (250 / 100)
    t = np.frombuffer(a.seg_host[1].numpy().tobytes(), dtype=SEG_DTYPE)
    assert np.isclose(t[0]["lr"], 0.01 * 0.1 ** 2.5, rtol=1e-5) and np.isclose(t[1]["lr"], 0.02)
    a.prepare_hyper(10 ** 9, 0)                               # clamped at lr:minimum_lr
    t = np.frombuffer(a.seg_host[0].numpy().tobytes(), dtype=SEG_DTYPE)
    assert np.isclose(t[0]["lr"], 1e-5)


def test_adam_table_carries_the_bias_corrected_rate():
    from cxxnet_b200.parallel.arena import SEG_DTYPE
    a, _ = _arena("adam")
    assert a.m2 is not None and a.kind == ops.UPD_ADAM
    a.prepare_hyper(4, 0)
    t = np.frombuffer(a.seg_host[0].numpy().tobytes(), dtype=SEG_DTYPE)
    b1, b2 = 0.1, 0.001
    want = 0.02 * np.sqrt(1 - (1 - b2) ** 5) / (1 - (1 - b1) ** 5)
    assert np.isclose(t[1]["adam_lr_t"], want, rtol=1e-5) and np.isclose(t[1]["beta1"], b1) and np.isclose(t[1]["beta2"], b2)


def test_collective_exchange_buckets_and_calls_on_a_mocked_device(monkeypatch):
    """parallel/dist_exchange.py without a GPU: CUDA streams / events replaced by stand-ins, a one-rank gloo group.
    Checks the bucket plan (backward order, small layers merged, trigger = lowest layer of the bucket), that a
    bucket's all-reduce covers exactly its chunk range and is followed by the update of that range, and the waits."""
    import torch.distributed as dist
    from cxxnet_b200.parallel import dist_exchange as dx
    from cxxnet_b200.parallel.arena import ParamArena

    class FakeEvent:
        def __init__(self, *a, **k): self.recorded = 0
        def record(self, *a): self.recorded += 1

    class FakeStream:
        def __init__(self, *a, **k): self.waits = 0
        def wait_event(self, ev): self.waits += 1

    class FakeCtx:
        def __init__(self, s): pass
        def __enter__(self): return self
        def __exit__(self, *a): return False
    main = FakeStream()
    monkeypatch.setattr(torch.cuda, "Event", FakeEvent)
    monkeypatch.setattr(torch.cuda, "Stream", FakeStream)
    monkeypatch.setattr(torch.cuda, "stream", FakeCtx)
    monkeypatch.setattr(torch.cuda, "current_stream", lambda *a, **k: main)
    monkeypatch.setattr(torch.cuda, "synchronize", lambda *a, **k: None)
    monkeypatch.setenv("CXN_XCHG_MIN_BUCKET_MB", "0.5")
    if not dist.is_initialized():
        dist.init_process_group("gloo", init_method="tcp://127.0.0.1:29636", rank=0, world_size=1)
    try:
        a = ParamArena(torch.device("cpu"), "sgd")
        a.upload_hyper = lambda slot=0: None
        sizes = {0: (96, 363), 4: (256, 1200), 8: (384, 2304), 16: (64, 4096)}      # layer index -> wmat shape
        for li, (r, c) in sizes.items():
            a.add(None, li, "wmat", r, c, c, _param("wmat"))
            a.add(None, li, "bias", 1, r, 0, _param("bias"))
        a.finalize()
        a.refresh_shadow = lambda: None
        updates = []
        a.update = lambda cb=0, ce=None: updates.append((cb, ce))

        class Net:
            pass
        net = Net()
        net.arena, net.rank, net.world_size, net.device = a, 0, 1, torch.device("cpu")
        ex = dx.CollectiveExchange(net)
        # backward order; a bucket closes once it holds >= 0.5 MB: [16] (1 MB), [8] (3.5 MB), [4] (1.2 MB), [0] (0.14 MB)
        assert [b["layers"] for b in ex.buckets] == [[16], [8], [4], [0]]
        assert all(b["trigger"] == b["layers"][-1] for b in ex.buckets)
        cover = sorted((b["cb"], b["ce"]) for b in ex.buckets)
        assert cover[0][0] == 0 and cover[-1][1] == a.nchunks and all(x[1] == y[0] for x, y in zip(cover, cover[1:]))
        a.g.fill_(1.0)
        ex.begin_step()
        ex.layer_done(12)                                   # a layer without weights: nothing happens
        assert updates == []
        for li in (16, 8, 4, 0):
            ex.layer_done(li)
        assert updates == [(b["cb"], b["ce"]) for b in ex.buckets]
        assert float(a.g.sum()) == float(a.g.numel())       # one rank: the sum over ranks is the gradient itself
        assert all(b["pending"] for b in ex.buckets)
        ex.wait_layer(8)
        assert [b["pending"] for b in ex.buckets] == [True, False, True, True]
        ex.wait_all()
        assert not any(b["pending"] for b in ex.buckets)
        assert ex.check_replicas() == 0.0
        ex.broadcast_weights()
        ex.publish_pending(); ex.after_replayed_step(); ex.gather_master()
    finally:
        dist.destroy_process_group()


def test_halo_tile_planning_takes_conv2_backward_data():
    """host side of conv_fprop_halo_kernel (cxn_conv_halo_plan, no GPU needed): AlexNet conv2's backward-data problem
    (dy 27x27x256 in two groups of 128, 5x5 flipped filter, pad 2, 48 outputs per group) is taken with the geometry the
    kernel's comments describe; problems it is not meant for are refused"""
    import ctypes
    L = ops.lib()
    L.cxn_conv_halo_plan.argtypes = [ctypes.c_int] * 15 + [ctypes.POINTER(ctypes.c_int)]
    L.cxn_conv_halo_plan.restype = ctypes.c_int
    out = (ctypes.c_int * 8)()
    # N, H, W, Cs, Cg, G, Ng, KH, KW, stride, pad_y, pad_x, OH, OW, halves
    assert L.cxn_conv_halo_plan(256, 27, 27, 256, 128, 2, 48, 5, 5, 1, 2, 2, 27, 27, 2, out) == 1
    wp, rows_half, tiles_img, patch_h, a_bytes, na, nb, smem = list(out)
    assert (wp, rows_half, tiles_img, patch_h) == (31, 4, 4, 12)
    assert a_bytes % 1024 == 0 and a_bytes >= ((4 + 4) * 31 + 4 + 128) * 128 and a_bytes >= 31 * 12 * 128
    assert na in (2, 3) and nb == 8 and smem <= 232448
    assert L.cxn_conv_halo_plan(256, 27, 27, 256, 128, 2, 48, 5, 5, 1, 2, 2, 27, 27, 1, out) == 1
    assert list(out)[:4] == [31, 4, 7, 8]                                      # single-half tiles: 4 rows each
    # refused: strided, channel groups that are not multiples of 64, more than 64 outputs, rows wider than a tile
    assert L.cxn_conv_halo_plan(8, 27, 27, 256, 128, 2, 48, 5, 5, 2, 2, 2, 14, 14, 2, out) == 0
    assert L.cxn_conv_halo_plan(8, 27, 27, 96, 48, 2, 128, 5, 5, 1, 2, 2, 27, 27, 2, out) == 0
    assert L.cxn_conv_halo_plan(8, 13, 13, 256, 256, 1, 384, 3, 3, 1, 1, 1, 13, 13, 2, out) == 0
    assert L.cxn_conv_halo_plan(8, 200, 200, 64, 64, 1, 32, 3, 3, 1, 1, 1, 200, 200, 2, out) == 0


def test_momentum_schedule_is_evaluated_once_per_update():
    """schedule_epoch ADDS to the momentum under momentum_schedule (reference param.h:76-94): building the arena
    must not count as an update, so the trajectory equals one evaluation per update like the ref backend's"""
    from cxxnet_b200.parallel.arena import ParamArena
    kv = dict(lr=0.01, momentum=0.5, momentum_schedule=1, base_momentum=0.5, final_momentum=0.9, saturation_epoch=10)
    a = ParamArena(torch.device("cpu"), "sgd")
    s = a.add(None, 0, "wmat", 4, 256, 256, _param("wmat", **kv))
    a.upload_hyper = lambda slot=0: None
    a.finalize()
    twin = _param("wmat", **kv)
    assert s.param.momentum == twin.momentum          # untouched by finalize()
    for epoch in range(4):
        a.prepare_hyper(epoch, epoch & 1)
        twin.schedule_epoch(epoch)
        assert s.param.momentum == twin.momentum
