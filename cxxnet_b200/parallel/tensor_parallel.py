"""Model-parallel (column-sharded) fully connected layers: ``fullc_shard = 1`` on several ranks.

The reference trains fully connected layers data parallel: every device holds all of W, pushes the whole
weight gradient and pulls the whole weight back every step (``src/updater/async_updater-inl.hpp:94-127``); its
``fullc_gather`` option (``:67-93,190-199``, ``src/layer/fullc_layer-inl.hpp:108-129``) reduces the traffic by
gathering inputs and output gradients.  On NVLink-connected GPUs the cheaper cut is the other one: rank ``r``
owns output columns ``[r*ns, (r+1)*ns)`` of the layer - the matching rows of W with their fp32 master, bf16
shadow and momentum - and W never moves.  What moves are activations:

forward of a sharded layer
    ``X_all`` (global batch x nin) is assembled on every rank by peer stores (``cxn_tp_push``): a row gather when
    the input node is data parallel (the first layer of a chain), a column gather when it is the output shard of
    the previous sharded layer.  One GEMM gives this rank's columns of the output for the WHOLE global batch.  The
    last layer of a chain hands its columns back to the rows' owners (all-to-all by peer stores), so the node
    after the chain is data parallel again.
backward
    the output gradient arrives as this rank's columns for the whole batch (all-to-all at the end of the chain,
    otherwise the result of the next layer's reduction).  Weight gradient and bias gradient are LOCAL GEMMs /
    column sums over the global batch - no gradient exchange, no weight publish, the updater runs on the owned
    rows only.  Backward-data of a column shard is a partial sum; every rank writes its bf16 partial into its own
    symmetric block and ``cxn_tp_reduce_pull`` sums the block each rank needs (its columns of the previous
    sharded layer, or its rows of the data-parallel input) out of all ranks' partials.

Between the layers of a chain the nodes are stored as ``[global batch][ns]`` (``Node.tp_rows``); ReLU (folded
into the GEMM epilogue), dropout and the other element-wise layers run on that layout unchanged.

Why no entry barriers (one flag barrier per collective, at its end): every buffer a peer writes into
(``X_all``, the last layer's output node and output-gradient shard) or reads from (the partials) is used once per
step.  A rank can only reach the next use of a buffer after it has passed a LATER barrier of the same step (at
least the reduction of the chain's first layer, or the first barrier of the next pass), and every rank issues
that barrier's signal after its own use of the buffer in stream order (weight gradient -> backward-data ->
signal).  Weight gradients run on a side stream; ``end_step`` joins it and closes every backward pass with one more
barrier, so the buffers those GEMMs read are free before any peer's next pass writes them.  The same argument covers
evaluation passes: all ranks issue the same sequence.
"""
from __future__ import annotations

import ctypes
from typing import Dict, List, Optional

import torch

from .. import ops

_ELEMENTWISE_OK = ("relu", "dropout", "sigmoid", "tanh", "xelu")


class TpLayer:
    __slots__ = ("layer_index", "layer", "nin", "nout", "ns", "c0", "my_cols", "first", "last", "ld_in",
                 "x_all", "x_all_off", "part", "part_off", "ys", "dys", "dys_off", "y_off", "y_ld_bytes", "y_elem",
                 "slot_in", "slot_out", "slot_dy", "slot_pull", "wseg", "bseg", "event", "pending", "prev", "next")


class TensorParallel:
    """plans, owns the buffers of, and runs the column-sharded fully connected layers of one net"""

    def __init__(self, net):
        self.net = net
        self.rank, self.world = net.rank, net.world_size
        self.layers: Dict[int, TpLayer] = {}
        self.chains: List[List[TpLayer]] = []
        self.shard_nodes = set()          # node indices stored as [global batch][ns]
        self.node_width: Dict[int, int] = {}
        self.counters = None
        self.off_flags = 0
        self.stream = None
        self._sigs = False

    # ------------------------------------------------------------------ planning (before buffers exist)
    def plan(self, consumers: Dict[int, int]) -> None:
        from ..layers import native
        net = self.net
        conns = net.connections
        chunk = ops.lib().cxn_update_chunk()
        producer_of = {}
        for i, c in enumerate(conns):
            if not c.layer.self_loop:
                for n in c.nodes_out:
                    producer_of[n.index] = i

        def shardable(i):
            c = conns[i]
            l = c.layer
            if not (c.native and isinstance(l, native.FullConnectLayer) and getattr(l, "fullc_shard", 0)
                    and c.type != 0 and not getattr(l, "fullc_gather", 0)
                    and id(l) not in getattr(net, "_shared_ids", ())):
                return False
            p = l.param
            ns = ops.round_up((p.num_hidden + self.world - 1) // self.world, 8)
            # shard boundaries on update chunks; 16-byte column blocks in every bf16 matrix that is pushed or pulled
            return (ns * p.num_input_node) % chunk == 0 and p.num_input_node % 8 == 0 and p.num_hidden % 8 == 0

        i = 0
        n = len(conns)
        while i < n:
            if not shardable(i):
                i += 1
                continue
            chain = [i]
            inner_nodes = []
            j = i
            while True:
                # walk from fc `j` over element-wise layers to the next shardable fc
                y = conns[j].nodes_out[0]
                k = j + 1
                nodes = [y]
                ok = True
                while k < n and conns[k].layer.type_name in _ELEMENTWISE_OK and conns[k].native \
                        and conns[k].nodes_in[0] in nodes and len(conns[k].nodes_in) == 1:
                    out = conns[k].nodes_out[0]
                    if out not in nodes:
                        nodes.append(out)
                    k += 1
                if not (k < n and shardable(k) and conns[k].nodes_in[0] in nodes):
                    break
                # every node in between must be private to the chain: each node is read only by the next layer
                # of the run (self loops do not count as consumers)
                for nd in nodes:
                    if consumers.get(nd.index, 0) != 1 or nd.loss_input:
                        ok = False
                if conns[k].layer.param.num_input_node != conns[j].layer.param.num_hidden:
                    ok = False
                if not ok:
                    break
                inner_nodes.extend(nodes)
                for nd in nodes:
                    self.node_width[nd.index] = ops.round_up(
                        (conns[j].layer.param.num_hidden + self.world - 1) // self.world, 8)
                chain.append(k)
                j = k
            tls = []
            for pos, li in enumerate(chain):
                l = conns[li].layer
                p = l.param
                t = TpLayer()
                t.layer_index, t.layer = li, l
                t.nin, t.nout = p.num_input_node, p.num_hidden
                t.ns = ops.round_up((t.nout + self.world - 1) // self.world, 8)
                t.c0 = min(self.rank * t.ns, t.nout)
                t.my_cols = max(0, min(t.ns, t.nout - t.c0))
                t.first, t.last = pos == 0, pos == len(chain) - 1
                t.prev = tls[-1] if tls else None
                t.next = None
                if tls:
                    tls[-1].next = t
                t.event, t.pending = None, False
                tls.append(t)
                self.layers[li] = t
                l.tp = t
            for nd in inner_nodes:
                self.shard_nodes.add(nd.index)
            self.chains.append(tls)
            i = chain[-1] + 1

    def active(self) -> bool:
        return bool(self.layers)

    def total_rows(self) -> int:
        return self.net.max_batch * self.world

    def shard_width(self, node_index: int) -> int:
        """columns of the [global batch][ns] storage of a chain-internal node"""
        return self.node_width[node_index]

    # ------------------------------------------------------------------ symmetric memory
    def symmetric_bytes(self) -> int:
        B = self.total_rows()
        total = ops.lib().cxn_tp_flag_bytes() + 4096
        for t in self.layers.values():
            ld_in = ops.round_up(t.nin, 8)
            total += 2 * (B * ld_in * 2 + 4096)               # X_all, partial
            if t.last:
                total += B * ops.round_up(t.ns, 8) * 2 + 4096    # output-gradient shard (pushed into by the peers)
                total += self.net.max_batch * ops.round_up(t.nout, 8) * 4 + 4096   # output node (bf16 or fp32)
        return total

    def finish(self, block, consumers) -> None:
        """carve the buffers out of the symmetric block and re-point the nodes the peers write into"""
        net = self.net
        L = ops.lib()
        B = self.total_rows()
        P, I, LL = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong
        L.cxn_tp_push.argtypes = [P, I, I, LL, I, P, P, LL, LL, P, I, P]
        L.cxn_tp_push.restype = I
        L.cxn_tp_reduce_pull.argtypes = [P, I, I, LL, I, P, LL, LL, LL, I, I, P, LL, I, P]
        L.cxn_tp_reduce_pull.restype = I
        self.block = block
        L.cxn_tp_set_multicast.argtypes = [P]
        L.cxn_tp_set_multicast(block.multicast_ptr or None)       # gathers / reductions through the switch when there is one
        self.off_flags, _ = block.carve(L.cxn_tp_flag_bytes())
        self.counters = torch.zeros(L.cxn_tp_counter_bytes() // 4, dtype=torch.int32, device=net.device)
        slot = 0
        for tls in self.chains:
            for t in tls:
                conn = net.connections[t.layer_index]
                x, y = conn.nodes_in[0], conn.nodes_out[0]
                t.ld_in = ops.round_up(t.nin, 8)
                t.x_all_off, xa = block.alloc(B * t.ld_in, torch.bfloat16)
                t.x_all = xa
                t.part_off, t.part = block.alloc(B * t.ld_in, torch.bfloat16)
                t.slot_in, t.slot_pull = slot, slot + 1
                slot += 2
                ns_pad = ops.round_up(t.ns, 8)
                if t.last:
                    t.slot_out, t.slot_dy = slot, slot + 1
                    slot += 2
                    t.dys_off, dys = block.alloc(B * ns_pad, torch.bfloat16)
                    t.dys = ops.DevMat(B, t.ns, ns_pad, tensor=dys.view(B, ns_pad))
                    # output columns of this rank for the whole batch, before they are handed to the rows' owners
                    if y.loss_input:
                        ld = y.f32.shape[1]
                        t.y_off, buf = block.alloc(net.max_batch * ld, torch.float32)
                        y.f32 = buf.view(net.max_batch, ld)
                        t.y_ld_bytes, t.y_elem = ld * 4, 4
                        t.ys = torch.zeros((B, ops.round_up(t.ns, 4)), dtype=torch.float32, device=net.device)
                    else:
                        ld = y.act.ld
                        rows = y.act.t.shape[0]
                        t.y_off, buf = block.alloc(rows * ld, torch.bfloat16)
                        new = ops.DevMat(y.act.rows, y.act.cols, ld, tensor=buf.view(rows, ld))
                        for nd in net.nodes:       # aliases of the node (folded ReLU) follow it
                            if nd.act is y.act:
                                nd.act = new
                        t.y_ld_bytes, t.y_elem = ld * 2, 2
                        t.ys = torch.zeros((B, ns_pad), dtype=torch.bfloat16, device=net.device)
                else:
                    t.ys = None       # the output node itself is stored [B][ns]
                    t.dys = None
                seg = net._seg_of
                t.wseg = seg[(t.layer_index, "wmat")]
                t.bseg = seg.get((t.layer_index, "bias"))
                t.event = torch.cuda.Event()
        if slot > L.cxn_tp_max_slots():
            raise ValueError("too many sharded fully connected layers")
        self.slot_end = slot
        slot += 1
        if slot > L.cxn_tp_max_slots():
            raise ValueError("too many sharded fully connected layers")
        self.stream = torch.cuda.Stream(device=net.device, priority=-1)
        self.ready = torch.cuda.Event()
        self.ready2 = torch.cuda.Event()
        self._open = False

    # ------------------------------------------------------------------ collectives
    def _push(self, t_slot: int, src_ptr: int, src_ld: int, dst_ld: int, desc: List[List[int]]) -> None:
        flat = (ctypes.c_longlong * (5 * self.world))()
        for r, d in enumerate(desc):
            for k in range(5):
                flat[5 * r + k] = int(d[k])
        ops.check(ops.lib().cxn_tp_push(self.block.peers_c, self.world, self.rank, self.off_flags, t_slot,
                                        self.counters.data_ptr(), src_ptr, src_ld, dst_ld, flat, 128, ops.stream()),
                  "cxn_tp_push")

    def _barrier(self, t_slot: int) -> None:
        """a push that moves nothing: signal, then wait for every rank's signal"""
        self._push(t_slot, self.counters.data_ptr(), 16, 16, [[0, 0, 0, 0, 0] for _ in range(self.world)])

    def _pull(self, t_slot: int, part_off: int, part_ld: int, pull_off: int, rows: int, row_bytes: int, dst_ptr: int,
              dst_ld: int) -> None:
        ops.check(ops.lib().cxn_tp_reduce_pull(self.block.peers_c, self.world, self.rank, self.off_flags, t_slot,
                                               self.counters.data_ptr(), part_off, part_ld, pull_off, rows, row_bytes,
                                               dst_ptr, dst_ld, 128, ops.stream()), "cxn_tp_reduce_pull")

    # ------------------------------------------------------------------ forward / backward of one sharded layer
    def fc_forward(self, layer, is_train: bool) -> None:
        net = self.net
        t: TpLayer = layer.tp
        conn = net.connections[t.layer_index]
        x, y = conn.nodes_in[0], conn.nodes_out[0]
        b, step, B, W = x.shape[0], net.max_batch, self.total_rows(), self.world
        ldb = t.ld_in * 2
        if t.first:
            # row gather: my rows (zero-filled up to `step`) into rows [rank*step, ...) of every rank's X_all
            desc = [[0, t.x_all_off + self.rank * step * ldb, step, b, t.nin * 2] for _ in range(W)]
            self._push(t.slot_in, x.act.ptr, x.act.ld * 2, ldb, desc)
        else:
            # column gather: my columns of the previous layer's output, all rows, into columns [c0, c0 + mine)
            pc0, pmine = t.prev.c0, t.prev.my_cols
            desc = [[0, t.x_all_off + pc0 * 2, B if pmine else 0, B, pmine * 2] for _ in range(W)]
            self._push(t.slot_in, x.act.ptr, x.act.ld * 2, ldb, desc)
        p = layer.param
        Wb = layer.bind["wmat"]
        bias = (layer.bind["bias"].w.data_ptr() + t.c0 * 4) if p.no_bias == 0 else None
        flags = ops.GEMM_RELU if layer.fuse_relu else 0
        if t.last:
            if t.y_elem == 4:
                dptr, ldd = t.ys.data_ptr(), t.ys.shape[1]
            else:
                dptr, ldd = t.ys.data_ptr(), t.ys.shape[1]
                flags |= ops.GEMM_OUT_BF16
        else:
            dptr, ldd = y.act.ptr, y.act.ld
            flags |= ops.GEMM_OUT_BF16
        if t.my_cols > 0:
            ops.gemm(t.x_all.data_ptr(), t.ld_in, 0, 0, Wb.shadow.data_ptr() + t.c0 * Wb.shadow_ld * 2, Wb.shadow_ld, 0, 0,
                     dptr, ldd, 0, bias, 0, B, t.my_cols, t.nin, flags=flags)
        if t.last:
            # all-to-all: rows [d*step, (d+1)*step) of my columns go to rank d's output node at column c0
            es = t.y_elem
            src_ld = t.ys.shape[1] * es
            desc = [[d * step * src_ld, t.y_off + t.c0 * es, step if t.my_cols else 0, step, t.my_cols * es]
                    for d in range(W)]
            self._push(t.slot_out, t.ys.data_ptr(), src_ld, t.y_ld_bytes, desc)

    def fc_backward(self, layer, prop_grad: bool) -> None:
        net = self.net
        t: TpLayer = layer.tp
        conn = net.connections[t.layer_index]
        x, y = conn.nodes_in[0], conn.nodes_out[0]
        b, step, B, W = x.shape[0], net.max_batch, self.total_rows(), self.world
        p = layer.param
        if t.last:
            # all-to-all: my rows of the output gradient, columns of rank d -> rows [rank*step, ...) of d's shard
            gy = y.grad
            ns_pad = t.dys.ld
            desc = []
            for d in range(W):
                dc0 = min(d * t.ns, t.nout)
                dcols = max(0, min(t.ns, t.nout - dc0))
                desc.append([dc0 * 2, t.dys_off + self.rank * step * ns_pad * 2, step if dcols else 0, b, dcols * 2])
            self._push(t.slot_dy, gy.ptr, gy.ld * 2, ns_pad * 2, desc)
            dys_ptr, dys_ld = t.dys.ptr, t.dys.ld
        else:
            dys_ptr, dys_ld = y.grad.ptr, y.grad.ld
        Wb = layer.bind["wmat"]
        a = net.arena
        acc = net.accumulate_grads
        main = torch.cuda.current_stream(net.device)
        # Weight gradient, bias gradient and the update of the owned rows run BESIDE the backward pass on the side
        # stream: only backward-data and the reduction are on the critical path.  The side stream starts when the
        # output-gradient shard is complete and updates the weights (which rewrites the bf16 shadow) only after the
        # backward-data GEMM has read them.  end_step() joins it before the step's closing barrier.
        self.ready.record(main)
        self.stream.wait_event(self.ready)
        if t.my_cols > 0:
            with torch.cuda.stream(self.stream):
                prev_lane = ops.lib().cxn_gemm_workspace_lane(1)     # split-K scratch of its own: runs concurrently with main
                try:
                    # gW[c0:c0+mine] (mine x nin) (+)= dYs^T . X_all  over the global batch
                    g_ptr = a.g.data_ptr() + (t.wseg.start + t.c0 * t.nin) * 4
                    ops.gemm(dys_ptr, dys_ld, 0, 1, t.x_all.data_ptr(), t.ld_in, 0, 1, g_ptr, t.nin, 0, None, 0,
                             t.my_cols, t.nin, B, flags=ops.GEMM_BETA1 if acc else 0)
                finally:
                    ops.lib().cxn_gemm_workspace_lane(max(prev_lane, 0))
                if t.bseg is not None:
                    ops.check(ops.lib().cxn_colsum_bf16(dys_ptr, B, t.my_cols, dys_ld,
                                                         a.g.data_ptr() + (t.bseg.start + t.c0) * 4, 1 if acc else 0,
                                                         ops.stream()), "colsum")
        need_pull = prop_grad
        if need_pull:
            if t.my_cols > 0:
                # partial dX (B x nin) = dYs (B x mine) . W[c0:c0+mine]
                ops.gemm(dys_ptr, dys_ld, 0, 0, Wb.shadow.data_ptr() + t.c0 * Wb.shadow_ld * 2, Wb.shadow_ld, 0, 1,
                         t.part.data_ptr(), t.ld_in, 0, None, 0, B, t.nin, t.my_cols, flags=ops.GEMM_OUT_BF16)
            else:
                t.part.zero_()
        self.ready2.record(main)
        with torch.cuda.stream(self.stream):
            if net.need_update:
                self.stream.wait_event(self.ready2)
                if t.my_cols > 0:
                    e0 = t.wseg.start + t.c0 * t.nin
                    a.update(e0 // a.chunk, (e0 + t.my_cols * t.nin + a.chunk - 1) // a.chunk)
                if t.bseg is not None:
                    # (foreign bias entries see a zero gradient; only the owner's entries are ever used or saved)
                    a.update(t.bseg.start // a.chunk, (t.bseg.start + t.bseg.n + a.chunk - 1) // a.chunk)
            t.event.record(self.stream)
            t.pending = True
        self._open = True
        if need_pull:
            ldb = t.ld_in * 2
            if t.first:
                # my rows, all columns -> data-parallel input gradient
                self._pull(t.slot_pull, t.part_off, ldb, self.rank * step * ldb, b, t.nin * 2, x.grad.ptr, x.grad.ld * 2)
            else:
                pc0, pmine = t.prev.c0, t.prev.my_cols
                self._pull(t.slot_pull, t.part_off, ldb, pc0 * 2, B if pmine else 0, pmine * 2, x.grad.ptr, x.grad.ld * 2)

    def end_step(self) -> None:
        """close a backward pass: join the side stream, then one barrier.  A peer's next forward pass overwrites this
        rank's X_all / output-gradient shards, which the side-stream weight-gradient GEMMs read; the barrier tells
        every peer that this rank is done with them (its signal is stream-ordered after the join)."""
        if not self._open:
            return
        self._open = False
        self.wait_all()
        self._barrier(self.slot_end)

    def wait_layer(self, layer_index: int) -> None:
        t = self.layers.get(layer_index)
        if t is not None and t.pending:
            torch.cuda.current_stream(self.net.device).wait_event(t.event)
            t.pending = False

    def wait_all(self) -> None:
        for li in self.layers:
            self.wait_layer(li)

    # ------------------------------------------------------------------ checkpoints / weight access
    def gather_master(self) -> None:
        """refresh the fp32 masters of the foreign column shards (rows of W, bias entries) from their owners"""
        import torch.distributed as dist
        from .symm import _RawCudaBuffer
        net = self.net
        self.wait_all()
        torch.cuda.synchronize(net.device)
        dist.barrier()
        a = net.arena
        off_w = net.symm_offsets["w"]
        for t in self.layers.values():
            for r in range(self.world):
                if r == self.rank:
                    continue
                c0 = min(r * t.ns, t.nout)
                cols = max(0, min(t.ns, t.nout - c0))
                if cols == 0:
                    continue
                pieces = [(t.wseg.start + c0 * t.nin, cols * t.nin)]
                if t.bseg is not None:
                    pieces.append((t.bseg.start + c0, cols))
                for e0, cnt in pieces:
                    holder = _RawCudaBuffer(self.block.peer_ptrs[r] + off_w + e0 * 4, cnt * 4)
                    peer = torch.as_tensor(holder, device=net.device).view(torch.float32)
                    a.w[e0:e0 + cnt].copy_(peer)
        torch.cuda.synchronize(net.device)
        dist.barrier()
