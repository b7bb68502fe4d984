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

import os

import torch

from .. import ops
from .symm import SymmetricBlock


class FcShard:
    """plan for one fully connected layer trained with ``fullc_gather = 1`` on several ranks."""
    __slots__ = ("layer_index", "layer", "rows_per_rank", "my_rows", "row0", "x_off", "gy_off", "x_all", "gy_all",
                 "ld_in", "ld_out", "wseg", "bseg", "slot_gather", "slot_pub", "event", "pending")


class GradExchange:
    """Per-step orchestration of the NVLink exchange.

    * ordinary weighted layers are grouped, in backward order, into buckets of at least
      ``min_bucket_bytes`` (small conv layers share one launch: at 8 GPUs a launch costs two
      ~20 us flag barriers regardless of size) and exchanged by ``cxn_fused_rs_update_ag``;
    * fully connected layers configured with ``fullc_gather = 1`` (reference
      ``async_updater-inl.hpp:67-93,115-126``) do not exchange their weight gradient at all: the
      ``[input | output-gradient]`` rows of every rank are all-gathered (a few MB instead of the
      151 MB fc6 gradient), each rank computes the weight gradient of ITS row shard of W over the
      whole global batch, updates that shard (momentum lives only there) and publishes the new
      bf16 weights to every rank.  fp32 masters of foreign shards are refreshed on demand
      (``gather_master``) for checkpoints / ``get_weight``.
    """

    def __init__(self, net, grid: int = 0):
        import os
        import torch.distributed as dist
        self.net = net
        self._fuse_gather = os.environ.get("CXN_FUSE_UPDATE_GATHER", "0") == "1"
        self.arena = net.arena
        self.rank, self.world = net.rank, net.world_size
        self.device = net.device
        self.grid = grid or int(os.environ.get("CXN_XCHG_GRID", "296"))
        # Bucket size: on four or more ranks every launch costs two all-rank flag barriers whose price is the skew
        # between the ranks, not bytes - measured on 8 GPUs (profiles/r2_scaling.md) one launch for all convolution
        # layers of AlexNet (9.3 MB, at the end of backward) beats three overlapped ones by 0.05 ms per step; on two
        # ranks small buckets overlapped with the backward pass stay ahead.
        default_mb = "4" if self.world < 4 else "64"
        self.min_bucket_bytes = int(float(os.environ.get("CXN_XCHG_MIN_BUCKET_MB", default_mb)) * (1 << 20))
        self.seq = 0
        L = ops.lib()
        P, I, LL, U = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_uint
        L.cxn_fused_rs_update_ag.argtypes = [P, I, I, LL, LL, LL, LL, LL, LL, P, P, LL, LL, I, U, I, P, I, P]
        L.cxn_fused_rs_update_ag.restype = I
        L.cxn_symm_broadcast.argtypes = [P, I, I, LL, LL, LL, I, U, I, P, I, P]
        L.cxn_symm_broadcast.restype = I
        L.cxn_symm_allgather.argtypes = [P, I, I, LL, P, LL, LL, LL, P, LL, LL, LL, I, U, I, P, I, P]
        L.cxn_symm_allgather.restype = I
        L.cxn_xchg_set_epoch_ptr.argtypes = [P]
        L.cxn_xchg_set_epoch_ptr.restype = I
        if self.world > L.cxn_xchg_max_ranks():
            raise ValueError("too many ranks for the fused exchange")
        self.block: SymmetricBlock = net.symm_block
        self.off = net.symm_offsets
        # NVSwitch multicast address of the block (parallel/symm.py, kind "nvls"): the fused kernel then reduces with
        # multimem.ld_reduce and publishes with multimem.st (process-wide setting: the newest exchange wins)
        L.cxn_xchg_set_multicast.argtypes = [P]
        L.cxn_xchg_set_multicast(self.block.multicast_ptr or None)
        self.counter = torch.zeros(4, dtype=torch.int32, device=self.device)
        self.counter_main = torch.zeros(4, dtype=torch.int32, device=self.device)
        a = self.arena
        self.shards: Dict[int, FcShard] = getattr(net, "fc_shards", {})
        # buckets over the non-sharded weighted layers, merged in backward order
        by_layer: Dict[int, List[int]] = {}
        tp_layers = net.tp.layers if getattr(net, "tp", None) is not None else {}
        for si, s in enumerate(a.segments):
            if s.layer_index not in self.shards and s.layer_index not in tp_layers:
                by_layer.setdefault(s.layer_index, []).append(si)
        self.buckets: List[Dict] = []
        cur = None
        for li in sorted(by_layer, reverse=True):
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
        for i, b in enumerate(self.buckets):
            b["slot"] = i
        nslots = len(self.buckets)
        for sh in self.shards.values():
            sh.slot_gather, sh.slot_pub = nslots, nslots + 1
            nslots += 2
            sh.event = torch.cuda.Event()
            sh.pending = False
        if nslots + 1 > L.cxn_xchg_max_slots():
            raise ValueError("too many weighted layers for the fused exchange")
        self.bucket_of_layer = {}
        for b in self.buckets:
            for li in b["layers"]:
                self.bucket_of_layer[li] = b
        self.stream = torch.cuda.Stream(device=self.device, priority=-1)
        self.ready = torch.cuda.Event()
        # Per-step part of every kernel's sequence number: a device word bumped on the main stream at the
        # start of each backward pass (begin_step).  The host part `self.seq` only moves outside steps, so
        # a step captured into a CUDA graph replays with advancing barriers.
        # timing experiments only (results are wrong): comma list of conv | gather | publish | fcgemm
        self._skip = set(filter(None, os.environ.get("CXN_XCHG_SKIP", "").split(",")))
        # The bf16 rows a rank owns are published to the other ranks at the START of the next forward pass
        # (on the exchange stream, under the convolution layers) instead of at the end of backward, where the
        # transfer competed with the convolution buckets for NVLink and was not hidden (8 GPUs: 181 us).
        # Measured: 8 GPUs 1.016 -> 0.957 ms/step, 4 GPUs 1.184 -> 1.149, but 2 GPUs 1.616 -> 1.687 (half of all
        # fully connected weights then cross one link while the first convolutions run), hence world >= 4.
        env = os.environ.get("CXN_XCHG_DEFER_PUBLISH")
        self.defer_publish = (env != "0") if env is not None else self.world >= 4
        self._pending_pub: List[FcShard] = []
        self.epoch = torch.zeros(1, dtype=torch.int64, device=self.device)
        L.cxn_xchg_set_epoch_ptr(self.epoch.data_ptr())
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
        # every exchange kernel of the previous step has been waited for by this step's forward pass (or
        # by the join at the end of a captured step), so nothing in flight reads the word while it changes
        ops.check(ops.lib().cxn_counter_add(self.epoch.data_ptr(), 1, ops.stream()), "counter_add")

    def layer_done(self, layer_index: int) -> None:
        """called right after layer ``layer_index`` finished its backward on the main stream."""
        b = self.bucket_of_layer.get(layer_index)
        if b is None or b["trigger"] != layer_index or "conv" in self._skip:
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
        sh = self.shards.get(layer_index)
        if sh is not None and sh.pending:
            torch.cuda.current_stream(self.device).wait_event(sh.event)
            sh.pending = False

    def wait_all(self) -> None:
        for b in self.buckets:
            self.wait_layer(b["trigger"])
        for li in self.shards:
            self.wait_layer(li)
        if getattr(self.net, "tp", None) is not None:
            self.net.tp.wait_all()

    # ------------------------------------------------------------------ K18: sharded fully connected layers
    def _allgather(self, pieces, slot, counter, stream) -> None:
        """pieces: list of (src_ptr, dst_off, slot_bytes, valid_bytes)"""
        p = list(pieces) + [(None, 0, 0, 0)] * (2 - len(pieces))
        ops.check(ops.lib().cxn_symm_allgather(
            self.block.peers_c, self.world, self.rank, self.off["flags"],
            p[0][0], p[0][1], p[0][2], p[0][3], p[1][0], p[1][1], p[1][2], p[1][3], len(pieces),
            self.seq, slot, counter.data_ptr(), 64, stream), "cxn_symm_allgather")

    def fc_backward(self, layer, need_update: bool) -> None:
        """weight-gradient path of a sharded fully connected layer; called from the layer's backward on
        the main stream, after its backward-data GEMM has been issued."""
        net, a = self.net, self.arena
        sh = self.shards[layer.layer_index]
        x, y = layer.nodes_in[0], layer.nodes_out[0]
        b = x.shape[0]
        step = net.max_batch
        main = torch.cuda.current_stream(self.device)
        total_b = step * self.world
        p = layer.param
        nin, nout = p.num_input_node, p.num_hidden
        # Everything runs on the exchange stream, overlapped with the rest of backward.  The stream waits
        # for this rank's backward-data GEMM; the gather's entry barrier therefore also tells every rank
        # that all of them are done reading this layer's weights, which the publish below overwrites.
        self.ready.record(main)
        self.stream.wait_event(self.ready)
        with torch.cuda.stream(self.stream):
            # 1. gather [x | gy] rows of every rank (zero-filled up to `step` rows per rank); neither buffer
            #    is written again before this layer's next forward, which waits for this stream
            if "gather" not in self._skip:
                self._allgather([(x.act.ptr, sh.x_off, step * sh.ld_in * 2, b * sh.ld_in * 2),
                                 (y.grad.ptr, sh.gy_off, step * sh.ld_out * 2, b * sh.ld_out * 2)],
                                sh.slot_gather, self.counter, self.stream.cuda_stream)
            acc = net.accumulate_grads
            # this step updates and nothing accumulates: the updater runs in the weight-gradient epilogue over the
            # owned rows (cxn_gemm_wgrad_update) - their gradient never reaches memory
            # (opt-in here, CXN_FUSE_UPDATE_GATHER=1: on two GPUs the step was slower with it, 1.511-1.517 ms against
            # 1.494 - this GEMM shares the SMs with the backward pass and the fused epilogue holds them longer)
            fuse = bool(need_update and not acc and net.fuse_update and self._fuse_gather and sh.my_rows > 0
                        and nin % 4 == 0 and "fcgemm" not in self._skip)
            if fuse:
                e0 = sh.wseg.start + sh.row0 * nin
                sld = sh.wseg.shadow_ld
                ops.check(ops.lib().cxn_gemm_wgrad_update(
                    sh.gy_all.data_ptr() + sh.row0 * 2, sh.ld_out, sh.x_all.data_ptr(), sh.ld_in, sh.my_rows, nin, total_b,
                    a.w.data_ptr() + 4 * e0, a.m.data_ptr() + 4 * e0, (a.m2.data_ptr() + 4 * e0) if a.m2 is not None else None,
                    a.shadow.data_ptr() + (sh.wseg.shadow_off + sh.row0 * sld) * 2, nin, sld, a.hyper_ptr(sh.wseg), a.kind,
                    ops.stream()), "cxn_gemm_wgrad_update")
            elif sh.my_rows > 0 and "fcgemm" not in self._skip:
                # gW[row0:row0+my_rows] = GY_all[:, shard]^T . X_all  over the global batch
                gy_ptr = sh.gy_all.data_ptr() + sh.row0 * 2
                g_ptr = a.g.data_ptr() + (sh.wseg.start + sh.row0 * nin) * 4
                # (split-K scratch lane 1: this GEMM runs concurrently with the main stream's backward GEMMs)
                prev_lane = ops.lib().cxn_gemm_workspace_lane(1)
                try:
                    ops.gemm(gy_ptr, sh.ld_out, 0, 1, sh.x_all.data_ptr(), sh.ld_in, 0, 1, g_ptr, nin, 0, None, 0,
                             sh.my_rows, nin, total_b, flags=ops.GEMM_BETA1 if acc else 0)
                finally:
                    ops.lib().cxn_gemm_workspace_lane(max(prev_lane, 0))
            if sh.bseg is not None:
                ops.check(ops.lib().cxn_colsum_bf16(sh.gy_all.data_ptr(), total_b, nout, sh.ld_out,
                                                     a.g.data_ptr() + sh.bseg.start * 4, 1 if acc else 0,
                                                     ops.stream()), "colsum")
            if need_update:
                if sh.my_rows > 0 and not fuse:
                    e0 = sh.wseg.start + sh.row0 * nin
                    a.update(e0 // a.chunk, (e0 + sh.my_rows * nin + a.chunk - 1) // a.chunk)
                if sh.bseg is not None:
                    a.update(sh.bseg.start // a.chunk, (sh.bseg.start + sh.bseg.n + a.chunk - 1) // a.chunk)
                # 3. publish the updated bf16 rows to every rank (now, or at the start of the next forward)
                if self.defer_publish:
                    if sh not in self._pending_pub:
                        self._pending_pub.append(sh)
                else:
                    self._publish(sh)
            sh.event.record(self.stream)
            sh.pending = True

    def _publish(self, sh) -> None:
        """all-gather of this rank's updated bf16 rows of one sharded layer (exchange stream)"""
        a = self.arena
        sld = sh.wseg.shadow_ld
        src = a.shadow.data_ptr() + (sh.wseg.shadow_off + sh.row0 * sld) * 2
        dst = self.off["shadow"] + sh.wseg.shadow_off * 2
        if "publish" not in self._skip:
            self._allgather([(src, dst, sh.rows_per_rank * sld * 2, sh.my_rows * sld * 2)],
                            sh.slot_pub, self.counter, self.stream.cuda_stream)

    def publish_pending(self) -> None:
        """called at the start of every forward pass: launch the deferred publishes on the exchange stream;
        the forward of each sharded layer waits for its own (wait_layer).  Every rank has long passed the
        gather barrier of the step that produced these rows, i.e. nobody still reads the old ones."""
        if not self._pending_pub:
            return
        main = torch.cuda.current_stream(self.device)
        self.ready.record(main)
        self.stream.wait_event(self.ready)
        with torch.cuda.stream(self.stream):
            for sh in self._pending_pub:
                self._publish(sh)
                sh.event.record(self.stream)
                sh.pending = True
        self._pending_pub = []

    def after_replayed_step(self) -> None:
        """a replayed CUDA graph ran the sharded updates without touching host state: their rows are
        unpublished until the next forward pass (replayed or eager) does it"""
        if self.defer_publish:
            self._pending_pub = [sh for sh in self.shards.values()]

    def gather_master(self) -> None:
        """refresh the fp32 masters of foreign weight shards from their owners (collective)."""
        self.wait_all()          # bucket kernels still pending on the exchange stream write arena.w
        if getattr(self.net, "tp", None) is not None:
            self.net.tp.gather_master()      # column-sharded layers: owners' rows of W and bias entries
        if not self.shards:
            return
        import torch.distributed as dist
        from .symm import _RawCudaBuffer
        torch.cuda.synchronize(self.device)
        dist.barrier()
        a = self.arena
        for sh in self.shards.values():
            nin = sh.layer.param.num_input_node
            nout = sh.layer.param.num_hidden
            for r in range(self.world):
                if r == self.rank:
                    continue
                row0 = r * sh.rows_per_rank
                rows = max(0, min(sh.rows_per_rank, nout - row0))
                if rows == 0:
                    continue
                e0 = sh.wseg.start + row0 * nin
                nbytes = rows * nin * 4
                holder = _RawCudaBuffer(self.block.peer_ptrs[r] + self.off["w"] + e0 * 4, nbytes)
                peer = torch.as_tensor(holder, device=self.device).view(torch.float32)
                a.w[e0:e0 + rows * nin].copy_(peer)
        torch.cuda.synchronize(self.device)
        dist.barrier()

    def check_replicas(self) -> float:
        """max |w_rank - w_rank0| over all weights (the reference's ``test_on_server`` idea)."""
        import torch.distributed as dist
        self.wait_all()
        self.gather_master()
        ref = self.arena.w.clone()
        dist.broadcast(ref, 0)
        d = (self.arena.w - ref).abs().max()
        dist.all_reduce(d, op=dist.ReduceOp.MAX)
        return float(d.item())
