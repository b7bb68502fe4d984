"""Flat parameter arenas for the b200 backend.

Every ``(layer, tag)`` weight of a net becomes one *segment* of three flat fp32 device
buffers - weights ``w``, gradients ``g``, optimiser state ``m`` (``m2`` for adam) - plus a
bf16 *shadow* buffer holding the GEMM-ready copy of matrix weights (rows padded to 8
elements for TMA).  Segments are laid out in layer order and aligned to the update chunk, so
that

* one kernel launch applies sgd / nag / adam to every weight (``cxn_update``, K15), and
* a layer bucket is a contiguous byte range - exactly what the fused
  reduce-scatter + update + all-gather kernel (K17, ``parallel/exchange.py``) exchanges over
  NVLink, in place of the reference's per-key parameter server
  (reference ``src/updater/async_updater-inl.hpp:94-127``, key = layer*4 + tag
  ``updater.h:150-173``).
"""
from __future__ import annotations

import math
from typing import Callable, Dict, List, Optional

import numpy as np
import torch

from .. import ops
from ..updaters import UpdaterParam

SEG_DTYPE = np.dtype([
    ("start", "<i8"), ("n", "<i8"), ("shadow_off", "<i8"), ("cols", "<i4"), ("shadow_ld", "<i4"),
    ("lr", "<f4"), ("wd", "<f4"), ("momentum", "<f4"), ("clip", "<f4"),
    ("adam_lr_t", "<f4"), ("beta1", "<f4"), ("beta2", "<f4"), ("flags", "<i4"),
])


class Segment:
    __slots__ = ("layer_index", "tag", "start", "n", "rows", "cols", "shadow_off", "shadow_ld", "param", "layer",
                 "beta1", "beta2", "nozero", "fused")


class ParamArena:
    def __init__(self, device: torch.device, updater_type: str,
                 alloc: Optional[Callable[[int, torch.dtype], torch.Tensor]] = None):
        self.device = device
        self.updater_type = updater_type
        self.kind = {"sgd": ops.UPD_SGD, "nag": ops.UPD_NAG, "adam": ops.UPD_ADAM}[updater_type]
        self.segments: List[Segment] = []
        self._alloc = alloc or (lambda n, dt: torch.zeros(n, dtype=dt, device=device))
        self.chunk = ops.lib().cxn_update_chunk()
        assert ops.lib().cxn_segment_bytes() == SEG_DTYPE.itemsize
        self.total = 0
        self.shadow_total = 0
        self.w = self.g = self.m = self.m2 = self.shadow = None
        self.seg_host = None
        self.seg_dev = None
        self.chunk_seg = None
        self.nchunks = 0
        self._hyper_arr = None
        self.fused_on = False   # this step's table tells cxn_update to skip the segments their producer updates

    # ------------------------------------------------------------------ layout
    def add(self, layer, layer_index: int, tag: str, rows: int, cols: int, shadow_ld: int, param: UpdaterParam,
            beta1: float = 0.1, beta2: float = 0.001) -> Segment:
        s = Segment()
        s.layer, s.layer_index, s.tag = layer, layer_index, tag
        s.rows, s.cols, s.n = rows, cols, rows * cols
        s.start = self.total
        self.total += ops.round_up(s.n, self.chunk)
        s.shadow_ld = shadow_ld
        if shadow_ld > 0:
            s.shadow_off = self.shadow_total
            self.shadow_total += ops.round_up(rows * shadow_ld, 128)
        else:
            s.shadow_off = -1
        s.param, s.beta1, s.beta2 = param, beta1, beta2
        s.nozero = False
        s.fused = False       # the layer's weight-gradient kernel can apply the update itself (cxn_gemm_wgrad_update)
        self.segments.append(s)
        return s

    def finalize(self) -> None:
        n = max(self.total, self.chunk)
        self.w = self._alloc(n, torch.float32)
        self.g = self._alloc(n, torch.float32)
        self.m = self._alloc(n, torch.float32)
        self.m2 = self._alloc(n, torch.float32) if self.kind == ops.UPD_ADAM else None
        self.shadow = self._alloc(max(self.shadow_total, 128), torch.bfloat16)
        self.nchunks = self.total // self.chunk
        cs = np.zeros(max(self.nchunks, 1), dtype=np.int32)
        for i, s in enumerate(self.segments):
            cs[s.start // self.chunk:(s.start + ops.round_up(s.n, self.chunk)) // self.chunk] = i
        self.chunk_seg = torch.from_numpy(cs).to(self.device)
        self._hyper_raw = [None, None]       # bytes last written to each pinned staging buffer
        self._hyper_copied = [None, None]    # events: the device has consumed the staging buffer
        self._hyper_slot = 0
        self.seg_host = [torch.zeros(max(len(self.segments), 1) * SEG_DTYPE.itemsize, dtype=torch.uint8) for _ in range(2)]
        if self.device.type == "cuda":
            self.seg_host = [t.pin_memory() for t in self.seg_host]
        self.seg_dev = torch.zeros_like(self.seg_host[0], device=self.device)
        # initial table: schedule_epoch is stateful under momentum_schedule (it ADDS to the momentum, as the
        # reference's ScheduleEpoch does), and this evaluation is not an update - put the momenta back afterwards
        saved = [s.param.momentum for s in self.segments]
        self.refresh_hyper(0)
        for s, mom in zip(self.segments, saved):
            s.param.momentum = mom

    # ------------------------------------------------------------------ views
    def w_view(self, s: Segment) -> torch.Tensor:
        return self.w[s.start:s.start + s.n]

    def g_view(self, s: Segment) -> torch.Tensor:
        return self.g[s.start:s.start + s.n]

    def hyper_ptr(self, s: Segment) -> int:
        """device address of ``{lr, wd, momentum, clip, adam_lr_t, beta1, beta2}`` of segment ``s`` in the per-step table"""
        i = self.segments.index(s)
        return self.seg_dev.data_ptr() + i * SEG_DTYPE.itemsize + SEG_DTYPE.fields["lr"][1]

    def shadow_view(self, s: Segment) -> Optional[torch.Tensor]:
        if s.shadow_off < 0:
            return None
        return self.shadow[s.shadow_off:s.shadow_off + s.rows * s.shadow_ld]

    def chunk_range(self, seg_begin: int, seg_end: int):
        """chunk interval covering segments [seg_begin, seg_end)."""
        a = self.segments[seg_begin].start // self.chunk
        last = self.segments[seg_end - 1]
        b = (last.start + ops.round_up(last.n, self.chunk)) // self.chunk
        return a, b

    # ------------------------------------------------------------------ per-step hyper parameters
    def prepare_hyper(self, epoch: int, slot: int = 0) -> bool:
        """Host half: evaluate every segment's schedule for update number ``epoch`` (reference
        ``src/updater/param.h:76-94`` ScheduleEpoch) and, if the table differs from what pinned
        staging buffer ``slot`` holds, rewrite that buffer.  Before rewriting, wait until the
        device has consumed its previous contents (the H2D copy is asynchronous, and a captured
        graph re-reads the same pinned address on every replay).  Two slots used alternately let
        the host run one step ahead of the device.  Returns whether the buffer was rewritten."""
        arr = self._hyper_arr
        if arr is None:
            # the fields that never change are filled once; a step only rewrites the scheduled columns
            arr = self._hyper_arr = np.zeros(max(len(self.segments), 1), dtype=SEG_DTYPE)
            for i, s in enumerate(self.segments):
                arr[i]["start"], arr[i]["n"], arr[i]["shadow_off"] = s.start, s.n, s.shadow_off
                arr[i]["cols"], arr[i]["shadow_ld"] = max(s.cols, 1), s.shadow_ld
                arr[i]["beta1"], arr[i]["beta2"] = (s.beta1, s.beta2) if self.kind == ops.UPD_ADAM else (0.0, 0.0)
        segs = self.segments
        if segs:
            for s in segs:
                s.param.schedule_epoch(epoch)
            n = len(segs)
            arr["lr"][:n] = [s.param.learning_rate for s in segs]
            arr["wd"][:n] = [s.param.wd for s in segs]
            arr["momentum"][:n] = [s.param.momentum for s in segs]
            arr["clip"][:n] = [s.param.clip_gradient for s in segs]
            # bit 0: the producing kernel overwrites this gradient; bit 1: it also applies this step's update
            fz = self.fused_on
            arr["flags"][:n] = [(1 if s.nozero else 0) | (2 if (fz and s.fused) else 0) for s in segs]
            if self.kind == ops.UPD_ADAM:
                arr["adam_lr_t"][:n] = [s.param.base_lr * math.sqrt(1.0 - (1.0 - s.beta2) ** (epoch + 1))
                                        / (1.0 - (1.0 - s.beta1) ** (epoch + 1)) for s in segs]
        raw = arr.tobytes()
        if raw == self._hyper_raw[slot]:
            return False
        if self._hyper_copied[slot] is not None:
            self._hyper_copied[slot].synchronize()
        self.seg_host[slot].numpy()[:] = np.frombuffer(raw, dtype=np.uint8)
        self._hyper_raw[slot] = raw
        return True

    def upload_hyper(self, slot: int = 0) -> None:
        """Device half (stream ordered, capturable): pinned table ``slot`` -> device table."""
        self.seg_dev.copy_(self.seg_host[slot], non_blocking=True)
        if not torch.cuda.is_current_stream_capturing():
            self.mark_hyper_consumed(slot)

    def mark_hyper_consumed(self, slot: int = 0) -> None:
        """the pinned table ``slot`` may be rewritten once the stream reaches this point"""
        if self._hyper_copied[slot] is None:
            self._hyper_copied[slot] = torch.cuda.Event()
        self._hyper_copied[slot].record()

    def refresh_hyper(self, epoch: int) -> None:
        self._hyper_slot ^= 1
        self.prepare_hyper(epoch, self._hyper_slot)
        self.upload_hyper(self._hyper_slot)

    # ------------------------------------------------------------------ kernels
    def update(self, chunk_begin: int = 0, chunk_end: Optional[int] = None) -> None:
        chunk_end = self.nchunks if chunk_end is None else chunk_end
        if chunk_end <= chunk_begin:
            return
        ops.check(ops.lib().cxn_update(
            self.w.data_ptr(), self.g.data_ptr(), self.m.data_ptr(),
            self.m2.data_ptr() if self.m2 is not None else None, self.shadow.data_ptr(),
            self.seg_dev.data_ptr(), self.chunk_seg.data_ptr(), chunk_begin, chunk_end,
            self.kind, ops.stream()), "cxn_update")

    def refresh_shadow(self) -> None:
        if self.nchunks:
            ops.check(ops.lib().cxn_refresh_shadow(self.w.data_ptr(), self.shadow.data_ptr(), self.seg_dev.data_ptr(),
                                                   self.chunk_seg.data_ptr(), self.nchunks, ops.stream()),
                      "cxn_refresh_shadow")
