"""Gradient exchange through ``torch.distributed`` collectives: the path for ranks that cannot map each
other's memory (``param_server = dist``: several nodes, or one node without peer access).

The reference's multi-node mode is an asynchronous ps-lite parameter server (reference
``Makefile:67-69``, ``src/nnet/nnet_ps_server.cpp:20-149``, ``example/MNIST/mpi.conf``): workers push
gradients to server processes that own the weights and the updater.  Here every rank keeps a full replica
and the exchange is synchronous, with the same placement as the NVLink path (``exchange.py``): when a
bucket's backward has finished, the exchange stream all-reduces (SUM) that contiguous range of the
gradient arena and runs the multi-tensor updater on it, while the main stream continues with the lower
layers; the next forward of those layers waits on the bucket's event (``UpdateWait``,
``async_updater-inl.hpp:140-143``).  Every rank applies the same update to the same sum, so replicas stay
bit-identical without a weight transfer and there is no server role (``update_on_server`` has nothing to
select).  Inside one NVLink domain the fused kernel of ``exchange.py`` is the fast path; this one is also
the "collective library" arm it is compared against (``CXN_EXCHANGE=collective``).
"""
from __future__ import annotations

import os
from typing import Dict, List

import torch

from .. import ops


class CollectiveExchange:
    """same interface as ``exchange.GradExchange`` (the executor and the trainer call either)."""

    def __init__(self, net):
        import torch.distributed as dist
        self.net = net
        self.arena = net.arena
        self.rank, self.world = net.rank, net.world_size
        self.device = net.device
        self.shards: Dict = {}                      # no row-sharded fully connected layers on this path
        self.min_bucket_bytes = int(float(os.environ.get("CXN_XCHG_MIN_BUCKET_MB", "4")) * (1 << 20))
        a = self.arena
        by_layer: Dict[int, List[int]] = {}
        for si, s in enumerate(a.segments):
            by_layer.setdefault(s.layer_index, []).append(si)
        self.buckets: List[Dict] = []
        cur = None
        for li in sorted(by_layer, reverse=True):       # backward order; arena ranges are contiguous per layer
            segs = by_layer[li]
            cb, ce = a.chunk_range(segs[0], segs[-1] + 1)
            if cur is not None and cur["cb"] == ce and (cur["ce"] - cur["cb"]) * a.chunk * 4 < self.min_bucket_bytes:
                cur["cb"] = cb
                cur["layers"].append(li)
                cur["trigger"] = li
            else:
                cur = {"layers": [li], "trigger": li, "cb": cb, "ce": ce, "event": torch.cuda.Event(),
                       "pending": False}
                self.buckets.append(cur)
        self.bucket_of_layer = {li: b for b in self.buckets for li in b["layers"]}
        self.stream = torch.cuda.Stream(device=self.device, priority=-1)
        self.ready = torch.cuda.Event()
        self._dist = dist
        dist.barrier()

    # ------------------------------------------------------------------ K19
    def broadcast_weights(self, root: int = 0) -> None:
        a = self.arena
        torch.cuda.synchronize(self.device)
        self._dist.broadcast(a.w, root)
        a.refresh_shadow()
        torch.cuda.synchronize(self.device)

    # ------------------------------------------------------------------ K17 through the collective library
    def begin_step(self) -> None:
        pass

    def layer_done(self, layer_index: int) -> None:
        b = self.bucket_of_layer.get(layer_index)
        if b is None or b["trigger"] != layer_index:
            return
        a = self.arena
        main = torch.cuda.current_stream(self.device)
        self.ready.record(main)
        self.stream.wait_event(self.ready)
        with torch.cuda.stream(self.stream):
            g = a.g[b["cb"] * a.chunk:b["ce"] * a.chunk]
            self._dist.all_reduce(g)                   # SUM: the loss layer already divides by the global batch
            a.update(b["cb"], b["ce"])                 # ops.stream() is the current (= exchange) stream
            b["event"].record(self.stream)
        b["pending"] = True

    def wait_layer(self, layer_index: int) -> None:
        b = self.bucket_of_layer.get(layer_index)
        if b is not None and b["pending"]:
            torch.cuda.current_stream(self.device).wait_event(b["event"])
            b["pending"] = False

    def wait_all(self) -> None:
        for b in self.buckets:
            self.wait_layer(b["trigger"])

    def publish_pending(self) -> None:
        pass

    def after_replayed_step(self) -> None:
        pass

    def gather_master(self) -> None:
        pass

    def fc_backward(self, layer, need_update: bool) -> None:      # pragma: no cover - never planned on this path
        raise RuntimeError("fullc_gather shards are not used with the collective exchange")

    def check_replicas(self) -> float:
        """max |w_rank - w_rank0| over all weights (``test_on_server``)."""
        self.wait_all()
        ref = self.arena.w.clone()
        self._dist.broadcast(ref, 0)
        d = (self.arena.w - ref).abs().max()
        self._dist.all_reduce(d, op=self._dist.ReduceOp.MAX)
        return float(d.item())


def ranks_share_one_host() -> bool:
    """True unless the launcher says the job spans machines.  torchrun exports LOCAL_WORLD_SIZE (ranks on
    this node) next to WORLD_SIZE; a launcher that sets neither is taken to be single-node.  No collective
    is issued, so this is safe before the rank has chosen its device."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_WORLD_SIZE", str(world)))
    return local >= world
