"""Multi-GPU gradient exchange (K17-K19 of SURVEY.md section 2.9).

Replaces the reference's ``AsyncUpdater`` <-> mshadow-ps push / pull protocol
(reference ``src/updater/async_updater-inl.hpp:94-143``) while keeping its observable semantics:

* per-(layer, tag) keys  -> per-layer *buckets* (contiguous chunk ranges of the parameter arena);
* SUM reduction over devices (the loss layer already divides by the global batch);
* the update for layer L is issued right after L's backward, lower priority for upper layers
  is implicit in the reverse-order issue (``priority = -layer_index``,
  ``updater_impl-inl.hpp:82``) - the buckets run on a high-priority side stream and overlap the
  rest of backward;
* ``UpdateWait`` before a layer's next forward -> the main stream waits on that bucket's event;
* initial weights are broadcast from rank 0 (the reference serialises device 0's model to the
  others, ``nnet_impl-inl.hpp:70-81``) with a peer-store kernel over the same symmetric block.

One launch of ``cxn_fused_rs_update_ag`` does reduce-scatter + updater + all-gather for a
bucket; there is no NCCL call and no separate update kernel on this path.
"""
from __future__ import annotations

import ctypes
from typing import Dict, List, Optional

import torch

from .. import ops
from .symm import SymmetricBlock


class GradExchange:
    def __init__(self, net, grid: int = 0):
        import torch.distributed as dist
        self.net = net
        self.arena = net.arena
        self.rank, self.world = net.rank, net.world_size
        self.device = net.device
        self.grid = grid or int(__import__('os').environ.get('CXN_XCHG_GRID', '296'))
        self.seq = 0
        L = ops.lib()
        P, I, LL, U = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_uint
        L.cxn_fused_rs_update_ag.argtypes = [P, I, I, LL, LL, LL, LL, LL, LL, P, P, LL, LL, I, U, I, P, I, P]
        L.cxn_fused_rs_update_ag.restype = I
        L.cxn_symm_broadcast.argtypes = [P, I, I, LL, LL, LL, I, U, I, P, I, P]
        L.cxn_symm_broadcast.restype = I
        if self.world > L.cxn_xchg_max_ranks():
            raise ValueError("too many ranks for the fused exchange")
        self.block: SymmetricBlock = net.symm_block
        self.off = net.symm_offsets
        self.counter = torch.zeros(4, dtype=torch.int32, device=self.device)
        # buckets: one per weighted layer, in layer order; slot = bucket index
        a = self.arena
        self.buckets: List[Dict] = []
        by_layer: Dict[int, List[int]] = {}
        for si, s in enumerate(a.segments):
            by_layer.setdefault(s.layer_index, []).append(si)
        for li in sorted(by_layer):
            segs = by_layer[li]
            cb, ce = a.chunk_range(segs[0], segs[-1] + 1)
            self.buckets.append({"layer": li, "cb": cb, "ce": ce, "slot": len(self.buckets),
                                 "event": torch.cuda.Event(), "pending": False})
        if len(self.buckets) + 1 > L.cxn_xchg_max_slots():
            raise ValueError("too many weighted layers for the fused exchange")
        self.bucket_of_layer = {b["layer"]: b for b in self.buckets}
        self.stream = torch.cuda.Stream(device=self.device, priority=-1)
        self.ready = torch.cuda.Event()
        dist.barrier()

    # ------------------------------------------------------------------ K19
    def broadcast_weights(self, root: int = 0) -> None:
        """replicate the root's weights (and bf16 shadows) into every rank's arena."""
        a = self.arena
        L = ops.lib()
        torch.cuda.synchronize(self.device)
        import torch.distributed as dist
        dist.barrier()
        self.seq += 1
        slot = L.cxn_xchg_max_slots() - 1
        for off, nbytes in ((self.off["w"], a.w.numel() * 4), (self.off["shadow"], a.shadow.numel() * 2)):
            nbytes = nbytes // 16 * 16
            ops.check(L.cxn_symm_broadcast(self.block.peers_c, self.world, self.rank, self.off["flags"], off, nbytes,
                                           root, self.seq, slot, self.counter.data_ptr(), self.grid, ops.stream()),
                      "cxn_symm_broadcast")
            self.seq += 1
        torch.cuda.synchronize(self.device)
        dist.barrier()

    # ------------------------------------------------------------------ K17
    def begin_step(self) -> None:
        self.seq += 1

    def layer_done(self, layer_index: int) -> None:
        """called right after layer ``layer_index`` finished its backward on the main stream."""
        b = self.bucket_of_layer.get(layer_index)
        if b is None:
            return
        a = self.arena
        main = torch.cuda.current_stream(self.device)
        self.ready.record(main)
        self.stream.wait_event(self.ready)
        ops.check(ops.lib().cxn_fused_rs_update_ag(
            self.block.peers_c, self.world, self.rank, self.off["w"], self.off["g"], self.off["m"],
            self.off.get("m2", 0), self.off["shadow"], self.off["flags"], a.seg_dev.data_ptr(),
            a.chunk_seg.data_ptr(), b["cb"], b["ce"], a.kind, self.seq, b["slot"], self.counter.data_ptr(),
            max(1, min(self.grid, (b["ce"] - b["cb"] + 3) // 4)), self.stream.cuda_stream), "cxn_fused_rs_update_ag")
        b["event"].record(self.stream)
        b["pending"] = True

    def wait_layer(self, layer_index: int) -> None:
        """``UpdateWait``: the next forward of this layer must see the updated weights."""
        b = self.bucket_of_layer.get(layer_index)
        if b is not None and b["pending"]:
            torch.cuda.current_stream(self.device).wait_event(b["event"])
            b["pending"] = False

    def wait_all(self) -> None:
        for b in self.buckets:
            self.wait_layer(b["layer"])

    def check_replicas(self) -> float:
        """max |w_rank - w_rank0| over all weights (the reference's ``test_on_server`` idea)."""
        import torch.distributed as dist
        self.wait_all()
        ref = self.arena.w.clone()
        dist.broadcast(ref, 0)
        d = (self.arena.w - ref).abs().max()
        dist.all_reduce(d, op=dist.ReduceOp.MAX)
        return float(d.item())
