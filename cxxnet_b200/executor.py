"""Per-device graph executor.

Functional parity with reference ``NeuralNet<xpu>`` (``src/nnet/neural_net-inl.hpp:23-297``): nodes,
connections, updaters; ``forward``, ``backprop``, model save / load, batch-size adjustment.

Backends
--------
``ref``   every layer runs its reference math on fp32 NCHW torch tensors (CPU or CUDA), with
          the reference's in-place node reuse; one ``Updater`` object per (layer, tag).
``b200``  NHWC bf16 nodes with separate gradient buffers, native sm_100a kernels for the hot
          layers, a layout bridge for the rest, all weights in a ``ParamArena`` updated by one
          fused multi-tensor kernel (or, multi-GPU, by the fused reduce-scatter + update +
          all-gather kernel of ``parallel/exchange.py``).  ReLU layers that directly follow a
          conv / fullc are folded into the GEMM epilogue.
"""
from __future__ import annotations

import io
from typing import BinaryIO, Dict, List, Optional

import numpy as np
import torch

from . import netconfig as nc
from . import ops
from .layers import LabelInfo, Layer, Node, create_layer
from .layers import native
from .layers.native import _ParamBinding
from .netconfig import NetConfig
from .updaters import UpdaterParam, create_updater


class Connection:
    __slots__ = ("type", "layer", "nodes_in", "nodes_out", "native")

    def __init__(self):
        self.type = 0
        self.layer: Optional[Layer] = None
        self.nodes_in: List[Node] = []
        self.nodes_out: List[Node] = []
        self.native = False


class NeuralNet:
    def __init__(self, cfg: NetConfig, max_batch: int, seed: int, device: torch.device, backend: str = "ref",
                 silent: int = 0, world_size: int = 1, rank: int = 0, arena_alloc=None,
                 exchange_kind: str = "nvlink"):
        self.cfg = cfg
        self.timer = None        # utils.timing.LayerTimer when `layer_timing = 1`
        self.max_batch = max_batch
        self.seed = seed
        self.device = device
        self.backend = backend
        self.silent = silent
        self.world_size, self.rank = world_size, rank
        # "nvlink": fused kernels over peer-mapped memory (parallel/exchange.py); "collective": torch.distributed
        # all-reduce + local update for ranks without peer access (parallel/dist_exchange.py)
        self.exchange_kind = exchange_kind
        self.rng = torch.Generator()
        self.rng.manual_seed(seed)
        self.label_info = LabelInfo(cfg.label_name_map)
        self.nodes: List[Node] = []
        self.connections: List[Connection] = []
        self.updaters: List[list] = []
        # b200 state
        self.arena = None
        self.arena_alloc = arena_alloc
        self.exchange = None
        self.tp = None           # parallel.tensor_parallel.TensorParallel when fully connected layers are column-sharded
        self._shared_ids = set() # ids of layer objects used by several connections (share[tag])
        self.wgrad_stream = None
        self.fuse_update = False
        self.fuse_now = False
        # step trace (`step_trace = 1` / CXN_STEP_TRACE=1): device timestamps between the layers of a step, also
        # inside a captured graph; read with trace_report()
        self.trace_buf = None
        self.trace_names = []
        self.need_update = True
        self.accumulate_grads = False
        # (per-channel means, scale) of 8-bit input batches: x = (pixel - mean[c]) * scale, applied where the batch
        # enters the net (DataBatch.norm, set by the batch iterator); None = pixels are used as they are
        self.input_norm = None
        self.num_sms = 148
        self.step_counter = None
        self.in_stage = None
        if backend == "b200":
            ops.lib()  # fail loudly if the kernel library is missing
            self.num_sms = torch.cuda.get_device_properties(device).multi_processor_count
            self.step_counter = torch.zeros(1, dtype=torch.int64, device=device)

    # ------------------------------------------------------------------ structure
    def _init_net(self) -> None:
        cfg = self.cfg
        self.nodes = [Node(cfg.node_names[i], i) for i in range(cfg.num_nodes)]
        c, h, w = cfg.input_shape
        self.nodes[0].shape = [self.max_batch, c, h, w]
        for i in range(cfg.extra_data_num):
            es = cfg.extra_shape
            self.nodes[i + 1].shape = [self.max_batch, es[i * 3], es[i * 3 + 1], es[i * 3 + 2]]
        self.connections = []
        for i, info in enumerate(cfg.layers):
            c_ = Connection()
            c_.type = info.type
            c_.nodes_in = [self.nodes[j] for j in info.nindex_in]
            c_.nodes_out = [self.nodes[j] for j in info.nindex_out]
            if info.type == nc.kSharedLayer:
                assert 0 <= info.primary_layer_index < len(self.connections), \
                    "shared layer primary_layer_index exceed bound"
                c_.layer = self.connections[info.primary_layer_index].layer
                if not c_.layer.allow_sharing:
                    raise ValueError("some layer you set shared do not allow sharing")
            else:
                c_.layer = create_layer(info.type)
                c_.layer.rng = self.rng
                c_.layer.label_info = self.label_info
                c_.layer.layer_index = i
                c_.layer.device = self.device
            self.connections.append(c_)

    def _config_connections(self) -> None:
        for i, c in enumerate(self.connections):
            if c.type == nc.kSharedLayer:
                continue
            for k, v in self.cfg.defcfg:
                c.layer.set_param(k, v)
            for k, v in self.cfg.layercfg[i]:
                c.layer.set_param(k, v)

    def _init_connections(self) -> None:
        for c in self.connections:
            c.layer.init_connection(c.nodes_in, c.nodes_out)
            if c.type != nc.kSharedLayer:
                c.layer.nodes_in, c.layer.nodes_out = c.nodes_in, c.nodes_out

    def init_model(self) -> None:
        self._init_net()
        self._config_connections()
        for i, c in enumerate(self.connections):
            if self.silent == 0:
                nm = self.cfg.layers[i].name
                print("Initializing layer: %s" % (nm if nm else str(i)))
        self._init_connections()
        for c in self.connections:
            if c.type != nc.kSharedLayer:
                c.layer.init_model()
                self._layer_to_device(c.layer)

    def _layer_to_device(self, layer: Layer) -> None:
        layer.device = self.device
        for tag, w, g in layer.weights():
            if w is not None:
                layer.rebind(tag, w.to(self.device), g.to(self.device))
        if hasattr(layer, "master"):
            self._layer_to_device(layer.master)
            self._layer_to_device(layer.slave)
        if getattr(layer, "wmat", None) is not None and not layer.weights():
            layer.wmat = layer.wmat.to(self.device)  # fixconn

    def save_model(self, fo: BinaryIO) -> None:
        self.sync_weights_from_device()
        for c in self.connections:
            if c.type != nc.kSharedLayer:
                c.layer.save_model(fo)

    def load_model(self, fi: BinaryIO, infer_shapes: bool = True) -> None:
        """``infer_shapes = False`` only reads every layer's stored parameters and weights: enough to copy layers by
        name (finetune) or to inspect a checkpoint, and possible without the configuration - a stored net does not
        carry the keys of its weightless layers (pooling windows ...), so its shapes cannot be inferred from the
        file alone (the reference's finetune path runs the shape inference on the stored net all the same,
        ``nnet_impl-inl.hpp:101-113``)."""
        self._init_net()
        self._config_connections()
        for c in self.connections:
            if c.type != nc.kSharedLayer:
                c.layer.load_model(fi)
                self._layer_to_device(c.layer)
        if infer_shapes:
            self._init_connections()

    def copy_layer(self, index: int, fi: BinaryIO) -> None:
        """finetune: load one layer's weights from a stream (reference ``neural_net-inl.hpp`` CopyLayer)."""
        self.sync_weights_from_device()
        layer = self.connections[index].layer
        old = dict((t, (w, g)) for t, w, g in layer.weights())
        layer.load_model(fi)
        for t, w, g in layer.weights():
            if t in old and old[t][0] is not None and old[t][0].shape == w.shape:
                old[t][0].copy_(w.to(old[t][0].device))
                layer.rebind(t, old[t][0], old[t][1])
            else:
                layer.rebind(t, w.to(self.device), g.to(self.device))
        self.sync_weights_to_device()

    # ------------------------------------------------------------------ updaters / nodes
    def init_updaters(self) -> None:
        cfg = self.cfg
        if self.backend == "b200":
            self._setup_b200()
            return
        self.updaters = []
        for i, c in enumerate(self.connections):
            out = []
            if c.type != nc.kSharedLayer:
                for tag, w, g in c.layer.weights():
                    u = create_updater(cfg.updater_type, w, g, tag)
                    for k, v in cfg.defcfg:
                        u.set_param(k, v)
                    for k, v in cfg.layercfg[i]:
                        u.set_param(k, v)
                    if self.silent:
                        u.param.silent = 1
                    u.init()
                    out.append(u)
            self.updaters.append(out)

    def init_nodes(self) -> None:
        for n in self.nodes:
            if self.silent == 0:
                print("node[%s].shape: %d,%d,%d,%d" % ((n.name,) + tuple(n.shape)))

    # ------------------------------------------------------------------ b200 setup
    def _setup_b200(self) -> None:
        from .parallel.arena import ParamArena
        cfg = self.cfg
        ops.set_device_index(self.device.index if self.device.index is not None else torch.cuda.current_device())
        consumers: Dict[int, int] = {}
        for c in self.connections:
            for n in c.nodes_in:
                if not (c.layer.self_loop and n in c.nodes_out):
                    consumers[n.index] = consumers.get(n.index, 0) + 1
        producer: Dict[int, Connection] = {}
        # Weight sharing (`share[tag]`, reference neural_net-inl.hpp:238-244): one layer object serves several
        # connections.  Fully connected and convolution layers (same shapes at every use, ordinary channel counts)
        # and the element-wise activations stay native: the executor binds the connection's nodes to the layer before
        # each call and makes the later uses of a backward pass ACCUMULATE into the one weight gradient; no epilogue /
        # mask fusion is planned around a shared layer (a fused flag would apply to every use).  Other shared layer
        # kinds keep per-layer state and run through the bridge.
        shared = set(id(c.layer) for c in self.connections if c.type == nc.kSharedLayer)
        self._shared_ids = shared

        def share_ok(layer):
            uses = [k for k in self.connections if k.layer is layer]
            first = uses[0]
            same = all([n.shape[1:] for n in k.nodes_in] == [n.shape[1:] for n in first.nodes_in] and
                       [n.shape[1:] for n in k.nodes_out] == [n.shape[1:] for n in first.nodes_out] for k in uses)
            if isinstance(layer, native.FullConnectLayer):
                return same
            if isinstance(layer, native.ConvolutionLayer):
                p = layer.param
                return same and (p.num_input_channel // p.num_group) % 8 == 0 and all(k.nodes_in[0].index != 0 for k in uses)
            return isinstance(layer, native._NativeAct)
        for c in self.connections:
            c.native = bool(c.layer.has_b200) and (id(c.layer) not in shared or share_ok(c.layer))
            if c.native and hasattr(c.layer, "b200_supported"):
                if id(c.layer) in shared:      # (the layer's own node binding is the primary's; shapes are the same at every use)
                    prim = next(k for k in self.connections if k.layer is c.layer)
                    c.layer.nodes_in, c.layer.nodes_out = prim.nodes_in, prim.nodes_out
                c.native = bool(c.layer.b200_supported())
            if not c.layer.self_loop:
                for n in c.nodes_out:
                    producer[n.index] = c
            if isinstance(c.layer, native._NativeLoss) and c.native:
                c.nodes_in[0].loss_input = True
        # kernel-path planning (may ask for a padded channel pitch on the input node)
        pitch: Dict[int, int] = {}
        for c in self.connections:
            if c.native and hasattr(c.layer, "b200_plan") and c.type != nc.kSharedLayer:
                x = c.nodes_in[0]
                want = c.layer.b200_plan(self, x.index == 0, consumers.get(x.index, 0) == 1)
                if isinstance(want, tuple) and want[0] == "s2d":      # space-to-depth input node, 64 channels per block
                    x.s2d, want = want[1], 64
                elif isinstance(want, tuple):        # (channel pitch, row pitch in pixels)
                    want, x.wpitch = want
                if want:
                    pitch[x.index] = want
        # column-sharded fully connected layers (several ranks over peer-mapped memory): planned before the node
        # buffers exist, because the nodes between two sharded layers are stored [global batch][own columns]
        peer_mapped = self.world_size > 1 and self.arena_alloc is None and self.exchange_kind == "nvlink"
        if peer_mapped:
            from .parallel.tensor_parallel import TensorParallel
            tp = TensorParallel(self)
            tp.plan(consumers)
            if tp.active():
                self.tp = tp
        # node buffers
        for n in self.nodes:
            b, ch, h, w = n.shape
            n.shape[0] = self.max_batch
            rows = self.max_batch if n.is_mat else self.max_batch * h * (n.wpitch or w)
            if n.wpitch:
                rows += 64     # zeroed slack: the last span of the last image row reads a few pixels past the image
            cols = w if n.is_mat else ch
            if n.s2d:
                rows, cols = self.max_batch * ((h + 3) // 4) * ((w + 3) // 4), 64
            ld = ops.round_up(cols, 8) if n.is_mat else pitch.get(n.index, cols)
            if self.tp is not None and n.index in self.tp.shard_nodes:
                n.tp_rows = rows = self.tp.total_rows()
                cols = self.tp.shard_width(n.index)
                ld = ops.round_up(cols, 8)
            n.act = ops.DevMat(rows, cols, ld, device=self.device)
            n.grad = ops.DevMat(rows, cols, ld, device=self.device)
            if n.loss_input:
                n.f32 = torch.zeros((self.max_batch, ops.round_up(cols, 4)), dtype=torch.float32, device=self.device)
        # relu folding / aliasing
        for c in self.connections:
            l = c.layer
            if c.native and isinstance(l, native._NativeAct) and l.alias_ok and id(l) not in shared:
                x, y = c.nodes_in[0], c.nodes_out[0]
                if x is y or consumers.get(x.index, 0) != 1 or y.loss_input or x.loss_input:
                    continue
                y.act, y.grad = x.act, x.grad
                y.alias_of = x
                p = producer.get(x.index)
                if (p is not None and p.native and isinstance(l, native.ReluLayer)
                        and isinstance(p.layer, (native.FullConnectLayer, native.ConvolutionLayer))
                        and id(p.layer) not in shared and consumers.get(x.index, 0) == 1):
                    p.layer.fuse_relu = True
                    l.fused = True
        # fold the backward mask of an aliased ReLU into the pooling layer that consumes it
        for c in self.connections:
            l = c.layer
            if c.native and isinstance(l, native._NativePool) and not isinstance(l, native.InsanityPoolingLayer) \
                    and id(l) not in shared:
                x = c.nodes_in[0]
                p = producer.get(x.index)
                if (p is not None and p.native and isinstance(p.layer, native.ReluLayer) and x.alias_of is not None
                        and consumers.get(x.index, 0) == 1 and x.grad is x.alias_of.grad):
                    l.mask_relu = True
                    p.layer.skip_bwd = True
                    # ... and if that ReLU was folded into a convolution with a bias, this pooling layer's input
                    # gradient is the convolution's output gradient: it accumulates the bias gradient as well
                    pc = producer.get(x.alias_of.index)
                    if (pc is not None and pc.native and isinstance(pc.layer, native.ConvolutionLayer)
                            and pc.type != nc.kSharedLayer and pc.layer.param.no_bias == 0 and p.layer.fused
                            and type(l) is native.MaxPoolingLayer and x.grad.ld == x.grad.cols
                            and __import__("os").environ.get("CXN_FUSE_BIAS_GRAD", "0") == "1"):
                        l.bias_of = pc.layer
                        pc.layer.bias_grad_fused = True
        # ... or into the backward-data kernel of the convolution that consumes it
        for c in self.connections:
            l = c.layer
            if c.native and isinstance(l, native.ConvolutionLayer) and l.igemm_dgrad and id(l) not in shared:
                x = c.nodes_in[0]
                p = producer.get(x.index)
                if (p is not None and p.native and isinstance(p.layer, native.ReluLayer) and x.alias_of is not None
                        and consumers.get(x.index, 0) == 1 and x.grad is x.alias_of.grad and not p.layer.skip_bwd
                        and x.act.ld == x.grad.ld):
                    l.mask_relu_in = True
                    p.layer.skip_bwd = True
        # parameter arena (in peer-mapped symmetric memory when several ranks train together)
        self.symm_block = None
        self.symm_offsets = {}
        self.arena = ParamArena(self.device, cfg.updater_type, alloc=self.arena_alloc)
        self._seg_of = {}
        for i, c in enumerate(self.connections):
            if c.type == nc.kSharedLayer:
                continue
            for tag, w, g in c.layer.weights():
                up = UpdaterParam(tag)
                b1, b2 = 0.1, 0.001
                for k, v in list(cfg.defcfg) + list(cfg.layercfg[i]):
                    up.set_param(k, v)
                    if k == "beta1": b1 = float(v)
                    if k == "beta2": b2 = float(v)
                shadow = c.layer.b200_shadow(tag) if (c.native and hasattr(c.layer, "b200_shadow")) else None
                if shadow is not None:
                    rows, cols, sld = shadow
                else:
                    rows, cols, sld = 1, w.numel(), 0
                s = self.arena.add(c.layer, i, tag, rows, cols, sld, up, b1, b2)
                # gradients that their producer overwrites on the first micro-step (GEMM with beta = 0, column
                # sums without accumulation) need no zeroing by the updater; convolution weight gradients are
                # accumulated with atomics and do
                s.nozero = bool(c.native and ((isinstance(c.layer, native.FullConnectLayer) and tag == "wmat") or
                                              (isinstance(c.layer, (native.FullConnectLayer, native.ConvolutionLayer))
                                               and tag == "bias")))
                self._seg_of[(i, tag)] = s
        if peer_mapped:
            self._plan_fc_shards()
            self._alloc_symmetric()
        self.arena.finalize()
        if peer_mapped:
            self._finish_fc_shards()
            if self.tp is not None:
                self.tp.finish(self.symm_block, consumers)
        self._bind_arena()
        # The weight-gradient GEMM of a fully connected layer applies the updater in its epilogue
        # (cxn_gemm_wgrad_update) and cxn_update skips those segments: the gradient never round-trips through HBM
        # (18 instead of 26 bytes per parameter).  AlexNet, one GPU: 2.580 vs 2.592 ms/step with the WIDE epilogue
        # (whole 512-byte row pieces per instruction); with 128-byte pieces it lost (2.617) - profiles/r2_experiments.md.
        # CXN_FUSE_UPDATE=0 keeps GEMM + multi-tensor update.
        import os
        self.fuse_update = os.environ.get("CXN_FUSE_UPDATE", "1") != "0"
        for (i, tag), s in self._seg_of.items():
            l = self.connections[i].layer
            s.fused = bool(self.connections[i].native and isinstance(l, native.FullConnectLayer) and tag == "wmat"
                           and id(l) not in shared and l.shard is None and l.tp is None
                           and l.param.num_input_node % 4 == 0)
        for c in self.connections:
            if c.native and c.type != nc.kSharedLayer:
                c.layer.b200_setup(self)
        c0, h0, w0 = cfg.input_shape
        # two fp32 staging slots filled on a copy stream, so the H2D transfer of batch t+1 overlaps
        # the compute of batch t
        self.in_stage = [torch.zeros((self.max_batch, c0, h0, w0), dtype=torch.float32, device=self.device)
                         for _ in range(2)]
        # bf16 batches (a data pipeline that ships half the bytes over PCIe) get their own pair of slots,
        # allocated when the first one arrives
        self.in_stage_bf16 = None
        self.in_stage_u8 = None       # 8-bit batches (`input_dtype = u8`): normalised on the device, see input_norm
        self.stage_dtype = [torch.float32, torch.float32]
        self.hyper_prepared = False
        self.hyper_slot = 0
        self.copy_stream = torch.cuda.Stream(device=self.device)
        # side stream for weight gradients that run beside a sub-wave backward-data kernel (layers/native.py)
        import os
        # (a captured step with more than one stream launches every node a little slower - measured on one GPU at batch
        # 32: the layers gain 10 us, the step loses 17 - so by default only where the step has side streams anyway)
        want = os.environ.get("CXN_OVERLAP_WGRAD", "auto")
        on = want == "1" or (want == "auto" and self.world_size > 1 and self.exchange_kind == "nvlink")
        self.wgrad_stream = torch.cuda.Stream(device=self.device) if on else None
        self.stage_free = [torch.cuda.Event(), torch.cuda.Event()]
        self.stage_ready = [torch.cuda.Event(), torch.cuda.Event()]
        self.stage_idx = 0
        self.extra_stage = [torch.zeros(tuple(self.nodes[i + 1].shape), dtype=torch.float32, device=self.device)
                            for i in range(cfg.extra_data_num)]
        if self.world_size > 1 and self.exchange_kind == "collective":
            from .parallel.dist_exchange import CollectiveExchange
            self.exchange = CollectiveExchange(self)
        elif self.world_size > 1 and self.exchange_kind != "server":
            from .parallel.exchange import GradExchange
            self.exchange = GradExchange(self)

    def _plan_fc_shards(self) -> None:
        """fully connected layers with ``fullc_gather = 1``: row-sharded weight gradient / update."""
        from .parallel.exchange import FcShard
        self.fc_shards = {}
        a = self.arena
        for i, c in enumerate(self.connections):
            l = c.layer
            if not (c.native and isinstance(l, native.FullConnectLayer) and l.fullc_gather
                    and c.type != nc.kSharedLayer and id(l) not in self._shared_ids) or l.tp is not None:
                continue
            p = l.param
            nin, nout = p.num_input_node, p.num_hidden
            rs = ops.round_up((nout + self.world_size - 1) // self.world_size, 8)   # TMA needs 16-byte aligned column offsets
            if (rs * nin) % a.chunk != 0:
                continue   # shard boundaries must fall on update chunks; otherwise the plain bucket path is used
            sh = FcShard()
            sh.layer_index, sh.layer = i, l
            sh.rows_per_rank = rs
            sh.row0 = min(self.rank * rs, nout)
            sh.my_rows = max(0, min(rs, nout - sh.row0))
            sh.ld_in, sh.ld_out = c.nodes_in[0].act.ld, c.nodes_out[0].grad.ld
            sh.wseg = self._seg_of[(i, "wmat")]
            sh.bseg = self._seg_of.get((i, "bias"))
            self.fc_shards[i] = sh
            l.shard = sh

    def _finish_fc_shards(self) -> None:
        total_rows = self.max_batch * self.world_size
        for sh in self.fc_shards.values():
            sh.x_off, x = self.symm_block.alloc(total_rows * sh.ld_in, torch.bfloat16)
            sh.gy_off, gy = self.symm_block.alloc(total_rows * sh.ld_out, torch.bfloat16)
            sh.x_all, sh.gy_all = x, gy

    def _alloc_symmetric(self) -> None:
        """lay the arenas out inside one CUDA-IPC block that every rank maps (parallel/symm.py)."""
        from .parallel.symm import SymmetricBlock
        a = self.arena
        n = max(a.total, a.chunk)
        nbuf = 4 if a.kind == ops.UPD_ADAM else 3
        flag_bytes = ops.lib().cxn_xchg_flag_bytes()
        total = nbuf * (n * 4 + 4096) + max(a.shadow_total, 128) * 2 + 4096 + flag_bytes + 65536
        for sh in getattr(self, "fc_shards", {}).values():
            total += self.max_batch * self.world_size * (sh.ld_in + sh.ld_out) * 2 + 4096
        if self.tp is not None:
            total += self.tp.symmetric_bytes()
        self.symm_block = SymmetricBlock(total, self.device)
        off, _ = self.symm_block.carve(flag_bytes)
        self.symm_offsets["flags"] = off
        names = ["w", "g", "m"] + (["m2"] if a.kind == ops.UPD_ADAM else []) + ["shadow"]
        it = iter(names)

        def alloc(count, dtype):
            o, t = self.symm_block.alloc(count, dtype)
            self.symm_offsets[next(it)] = o
            return t
        a._alloc = alloc

    def _bind_arena(self) -> None:
        """copy layer weights (reference layout) into the arena and hand views back to the layers."""
        a = self.arena
        for (i, tag), s in self._seg_of.items():
            c = self.connections[i]
            layer = c.layer
            w_ref = dict((t, w) for t, w, _ in layer.weights())[tag]
            wv, gv = a.w_view(s), a.g_view(s)
            if c.native:
                wv.copy_(layer.weight_to_internal(tag, w_ref.to(self.device)).reshape(-1))
                layer.bind[tag] = _ParamBinding(wv, gv, a.shadow_view(s), s.shadow_ld, s.rows, s.cols, s)
            else:
                if w_ref.data_ptr() != wv.data_ptr():
                    wv.copy_(w_ref.reshape(-1))
                layer.rebind(tag, wv.view(w_ref.shape), gv.view(w_ref.shape))
        a.refresh_shadow()

    def sync_weights_from_device(self) -> None:
        """arena (internal layout) -> layer tensors (checkpoint layout) for native layers.
        With sharded fully connected layers this is a collective over all ranks."""
        if self.arena is None:
            return
        if self.exchange is not None:
            self.exchange.gather_master()      # (includes the column-sharded layers of self.tp)
        for (i, tag), s in self._seg_of.items():
            c = self.connections[i]
            if not c.native:
                continue
            layer = c.layer
            w_ref = dict((t, w) for t, w, _ in layer.weights())[tag]
            internal = self.arena.w_view(s)
            w_ref.copy_(layer.weight_from_internal(tag, internal).reshape(w_ref.shape))

    def sync_weights_to_device(self) -> None:
        if self.arena is not None:
            self._bind_arena()

    # ------------------------------------------------------------------ layout bridge (b200)
    def _node_load(self, n: Node, grad: bool = False) -> torch.Tensor:
        if n.tp_rows:
            raise ValueError("node %s lies between two column-sharded fully connected layers (fullc_shard = 1): every "
                             "rank holds only its columns of it; extract it from a run without fullc_shard" % n.name)
        b, c, h, w = n.shape
        if n.loss_input and not grad:
            return n.f32[:b, :w].reshape(b, 1, 1, w).clone()
        m = n.grad if grad else n.act
        if n.is_mat:
            return m.t[:b, :w].float().reshape(b, 1, 1, w)
        if n.s2d:       # [b][Hs][Ws][dy][dx][4] -> [b][c][4 Hs][4 Ws], cropped
            hs, ws = (h + 3) // 4, (w + 3) // 4
            t = m.t[:b * hs * ws, :64].reshape(b, hs, ws, 4, 4, 4).permute(0, 5, 1, 3, 2, 4).reshape(b, 4, 4 * hs, 4 * ws)
            return t[:, :c, :h, :w].float().contiguous()
        wp = n.wpitch or w
        return m.t[:b * h * wp, :c].reshape(b, h, wp, c)[:, :, :w].permute(0, 3, 1, 2).float().contiguous()

    def _node_store(self, n: Node, t: torch.Tensor, grad: bool = False) -> None:
        b, c, h, w = n.shape
        if n.loss_input and not grad:
            n.f32[:b, :w].copy_(t.reshape(b, w))
            return
        m = n.grad if grad else n.act
        if n.is_mat:
            m.t[:b, :w].copy_(t.reshape(b, w))
        elif n.s2d:
            hs, ws = (h + 3) // 4, (w + 3) // 4
            full = torch.zeros((b, 4, 4 * hs, 4 * ws), dtype=t.dtype, device=t.device)
            full[:, :c, :h, :w] = t
            m.t[:b * hs * ws, :64].copy_(full.reshape(b, 4, hs, 4, ws, 4).permute(0, 2, 4, 3, 5, 1).reshape(b * hs * ws, 64))
        else:
            wp = n.wpitch or w
            m.t[:b * h * wp].reshape(b, h, wp, -1)[:, :, :w, :c].copy_(t.permute(0, 2, 3, 1))

    def _bridge_forward(self, c: Connection, is_train: bool) -> None:
        for n in c.nodes_in:
            n.data = self._node_load(n)
        c.layer.forward(is_train, c.nodes_in, c.nodes_out)
        for n in set(c.nodes_in + c.nodes_out):
            self._node_store(n, n.data)
            n.data = None

    def _bridge_backprop(self, c: Connection, prop_grad: bool) -> None:
        if c.layer.self_loop:
            for n in c.nodes_in:
                n.data = self._node_load(n, grad=True)
        else:
            for n in c.nodes_in:
                n.data = self._node_load(n)
            for n in c.nodes_out:
                n.data = self._node_load(n, grad=True)
        c.layer.backprop(prop_grad, c.nodes_in, c.nodes_out)
        if prop_grad or c.layer.self_loop:
            for n in c.nodes_in:
                self._node_store(n, n.data, grad=True)
        for n in set(c.nodes_in + c.nodes_out):
            n.data = None

    # ------------------------------------------------------------------ execution
    def adjust_batch_size(self, b: int) -> None:
        assert self.max_batch >= b, "cannot set batch size larger than max batch"
        if b != self.nodes[0].shape[0]:
            for n in self.nodes:
                n.shape[0] = b
            for c in self.connections:
                c.layer.on_batch_size_changed(c.nodes_in, c.nodes_out)

    def _norm(self, channels: int):
        """(means per channel, scale) for an 8-bit batch"""
        if self.input_norm is None:
            return [0.0] * channels, 1.0
        mean, scale = self.input_norm
        mean = [float(v) for v in mean][:channels]
        return mean + [0.0] * (channels - len(mean)), float(scale)

    def set_input(self, batch: torch.Tensor, extra: Optional[List[torch.Tensor]] = None) -> None:
        """copy one (already device-sliced) host batch into node 0 (+ extra data nodes)."""
        b = batch.shape[0]
        self.adjust_batch_size(b)
        extra = extra or []
        if self.backend == "ref":
            x = batch.to(self.device, non_blocking=True).float().reshape(self.nodes[0].shape)
            if batch.dtype == torch.uint8:
                mean, scale = self._norm(x.shape[1])
                x = (x - torch.tensor(mean, dtype=torch.float32, device=self.device).view(1, -1, 1, 1)) * scale
            self.nodes[0].data = x
            for i, e in enumerate(extra):
                self.nodes[i + 1].data = e.to(self.device, non_blocking=True).float().reshape(self.nodes[i + 1].shape)
            return
        slot = self.stage_input(batch)
        self.convert_staged(slot)
        self.release_stage(slot)
        for i, e in enumerate(extra):
            self._stage_in(self.nodes[i + 1], self.extra_stage[i], e)

    # The three parts of set_input for node 0, separately callable: a training step captured into a
    # CUDA graph contains only the middle one (the staging buffers are persistent, one graph per slot).
    def stage_input(self, batch: torch.Tensor) -> int:
        """copy the batch into the next staging slot (pinned host memory goes through the copy stream,
        overlapping the step still running) and make the main stream wait for it; returns the slot"""
        b = batch.shape[0]
        self.stage_idx ^= 1
        slot = self.stage_idx
        main = torch.cuda.current_stream(self.device)
        if batch.dtype == torch.bfloat16:
            if self.in_stage_bf16 is None:
                self.in_stage_bf16 = [torch.zeros_like(t, dtype=torch.bfloat16) for t in self.in_stage]
            st = self.in_stage_bf16[slot][:b]
        elif batch.dtype == torch.uint8:
            if self.in_stage_u8 is None:
                self.in_stage_u8 = [torch.zeros_like(t, dtype=torch.uint8) for t in self.in_stage]
            st = self.in_stage_u8[slot][:b]
        else:
            st = self.in_stage[slot][:b]
        self.stage_dtype[slot] = st.dtype
        if batch.device.type == "cpu" and batch.is_pinned():
            self.copy_stream.wait_event(self.stage_free[slot])
            with torch.cuda.stream(self.copy_stream):
                st.copy_(batch.reshape(st.shape), non_blocking=True)
                self.stage_ready[slot].record(self.copy_stream)
            main.wait_event(self.stage_ready[slot])
        else:
            st.copy_(batch.reshape(st.shape), non_blocking=True)
        return slot

    def convert_staged(self, slot: int) -> None:
        """staging slot (fp32 NCHW) -> node 0 in the backend's layout"""
        n = self.nodes[0]
        b = n.shape[0]
        if self.stage_dtype[slot] == torch.uint8:
            src = self.in_stage_u8[slot][:b]
            _, c, h, w = n.shape
            mean, scale = self._norm(c)
            if n.s2d:
                m = list(mean) + [0.0] * (4 - len(mean))
                ops.check(ops.lib().cxn_nchw_to_s2d4(src.data_ptr(), 2, n.act.ptr, b, c, h, w, m[0], m[1], m[2], m[3], scale,
                                                     ops.stream()), "nchw_to_s2d4")
                return
            if n.wpitch and not n.is_mat and hasattr(ops.lib(), "cxn_nchw_u8_to_nhwc4_pitched"):
                m = list(mean) + [0.0] * (4 - len(mean))
                ops.check(ops.lib().cxn_nchw_u8_to_nhwc4_pitched(src.data_ptr(), n.act.ptr, b, c, h, w, n.wpitch, m[0], m[1],
                                                                 m[2], m[3], scale, ops.stream()),
                          "nchw_u8_to_nhwc4_pitched")
                return
            # other layouts: widen and normalise on the device, then the fp32 path
            f = self.in_stage[slot][:b]
            f.copy_(src)
            key = (tuple(mean), scale)
            if getattr(self, "_norm_key", None) != key:       # device copy of the means, made once (not under capture)
                self._norm_key = key
                self._norm_mean = torch.tensor(mean, dtype=torch.float32, device=self.device).view(1, c, 1, 1)
            f.sub_(self._norm_mean).mul_(scale)
            self._convert_in(n, f)
            return
        if self.stage_dtype[slot] == torch.bfloat16:
            src = self.in_stage_bf16[slot][:b]
            if n.s2d:
                _, c, h, w = n.shape
                ops.check(ops.lib().cxn_nchw_to_s2d4(src.data_ptr(), 1, n.act.ptr, b, c, h, w, 0.0, 0.0, 0.0, 0.0, 1.0,
                                                     ops.stream()), "nchw_to_s2d4")
                return
            if n.wpitch and not n.is_mat:
                _, c, h, w = n.shape
                ops.check(ops.lib().cxn_nchw_bf16_to_nhwc4_pitched(src.data_ptr(), n.act.ptr, b, c, h, w, n.wpitch,
                                                                   ops.stream()), "nchw_bf16_to_nhwc4_pitched")
                return
            self.in_stage[slot][:b].copy_(src)      # other layouts: widen on the device, then the fp32 path
        self._convert_in(n, self.in_stage[slot][:b])

    def release_stage(self, slot: int) -> None:
        self.stage_free[slot].record(torch.cuda.current_stream(self.device))

    def _stage_in(self, n: Node, stage: torch.Tensor, batch: torch.Tensor) -> None:
        b = n.shape[0]
        st = stage[:b]
        st.copy_(batch.reshape(st.shape), non_blocking=True)
        self._convert_in(n, st)

    def _convert_in(self, n: Node, st: torch.Tensor) -> None:
        b, c, h, w = n.shape
        L = ops.lib()
        if n.is_mat:
            ops.check(L.cxn_copy2d(st.data_ptr(), w, n.act.ptr, n.act.ld, b, w, 1, 0, ops.stream()), "copy2d")
        elif n.s2d:
            ops.check(L.cxn_nchw_to_s2d4(st.data_ptr(), 0, n.act.ptr, b, c, h, w, 0.0, 0.0, 0.0, 0.0, 1.0, ops.stream()),
                      "nchw_to_s2d4")
        elif n.wpitch:
            ops.check(L.cxn_nchw_f32_to_nhwc4_pitched(st.data_ptr(), n.act.ptr, b, c, h, w, n.wpitch, ops.stream()),
                      "nchw_to_nhwc4_pitched")
        else:
            ops.check(L.cxn_nchw_f32_to_nhwc_bf16(st.data_ptr(), n.act.ptr, b, c, h * w, n.act.ld, ops.stream()),
                      "nchw_to_nhwc")

    # ------------------------------------------------------------------ step trace
    def enable_trace(self) -> None:
        n = len(self.connections)
        self.trace_names = ["begin"] + ["f%d:%s" % (i, c.layer.type_name) for i, c in enumerate(self.connections)] + \
                           ["metric"] + ["b%d:%s" % (i, self.connections[i].layer.type_name) for i in range(n - 1, -1, -1)] + ["end"]
        self.trace_buf = torch.zeros(len(self.trace_names), dtype=torch.int64, device=self.device)

    def _mark(self, slot: int) -> None:
        if self.trace_buf is not None:
            ops.check(ops.lib().cxn_timestamp(self.trace_buf.data_ptr(), slot, ops.stream()), "timestamp")

    def trace_report(self) -> str:
        """time between consecutive marks of the LAST executed step (us): a layer's entry is the stream time from the
        previous mark to the mark after the layer's last launch, barrier waits included"""
        if self.trace_buf is None:
            return ""
        t = self.trace_buf.cpu().tolist()
        lines = ["step trace (us, device %globaltimer; last step)"]
        for k in range(1, len(t)):
            if t[k] and t[k - 1]:
                lines.append("%-24s %9.1f" % (self.trace_names[k], (t[k] - t[k - 1]) / 1e3))
        if t[0] and t[-1]:
            lines.append("%-24s %9.1f" % ("whole step", (t[-1] - t[0]) / 1e3))
        return "\n".join(lines)

    def forward(self, is_train: bool) -> None:
        tm = self.timer
        if self.backend == "ref":
            for i, c in enumerate(self.connections):
                t0 = tm.begin() if tm else None
                c.layer.forward(is_train, c.nodes_in, c.nodes_out)
                if tm:
                    tm.end("f", i, t0)
            return
        if self.exchange is not None:
            self.exchange.publish_pending()
        if is_train:
            self._mark(0)
        for i, c in enumerate(self.connections):
            if self.exchange is not None:
                self.exchange.wait_layer(i)   # UpdateWait: this layer's weights from the last exchange
            if self.tp is not None:
                self.tp.wait_layer(i)
            t0 = tm.begin() if tm else None
            if c.native:
                if id(c.layer) in self._shared_ids:       # a shared layer works on this connection's nodes
                    c.layer.nodes_in, c.layer.nodes_out = c.nodes_in, c.nodes_out
                c.layer.b200_forward(self, is_train)
            else:
                self._bridge_forward(c, is_train)
            if tm:
                tm.end("f", i, t0)
            if is_train:
                self._mark(1 + i)

    def backprop(self, prop_to_input: bool, need_update: bool, epoch: int) -> None:
        n = len(self.connections)
        tm = self.timer
        if self.backend == "ref":
            for i in range(n - 1, -1, -1):
                c = self.connections[i]
                t0 = tm.begin() if tm else None
                c.layer.backprop(i != 0 or prop_to_input, c.nodes_in, c.nodes_out)
                if tm:
                    tm.end("b", i, t0)
                if need_update and self.exchange is not None:
                    # parameter-server run (parallel/ps_server.py): the gradient goes to the server, which updates
                    self.exchange.layer_done(i)
                elif need_update:
                    for u in self.updaters[i]:
                        if self.world_size > 1:
                            # reference-backend data parallelism: plain sum over ranks (gloo / nccl)
                            import torch.distributed as dist
                            dist.all_reduce(u.dw)
                        u.update(epoch)
            return
        self.need_update = need_update
        self.fuse_now = False
        if need_update:
            fuse = self.fuse_update_ok() and not self.accumulate_grads
            if self.hyper_prepared:          # the host half already ran (captured step)
                fuse = fuse and self.arena.fused_on
                self.arena.upload_hyper(self.hyper_slot)
            else:
                self.arena.fused_on = fuse
                self.arena.refresh_hyper(epoch)
            self.fuse_now = fuse
        if self.exchange is not None:
            self.exchange.begin_step()
        self._mark(1 + n)          # the train metric and the hyper-parameter upload sit between forward and backward
        base_acc = self.accumulate_grads
        seen = set()
        for i in range(n - 1, -1, -1):
            c = self.connections[i]
            prop = i != 0 or prop_to_input
            t0 = tm.begin() if tm else None
            if c.native:
                if id(c.layer) in self._shared_ids:
                    # every use after the first one of this pass adds to the shared weight gradient
                    c.layer.nodes_in, c.layer.nodes_out = c.nodes_in, c.nodes_out
                    self.accumulate_grads = base_acc or id(c.layer) in seen
                    seen.add(id(c.layer))
                c.layer.b200_backprop(self, prop)
                self.accumulate_grads = base_acc
            else:
                self._bridge_backprop(c, prop)
            if tm:
                tm.end("b", i, t0)
            if need_update and self.exchange is not None:
                self.exchange.layer_done(i)
            self._mark(2 + n + (n - 1 - i))
        if need_update and self.exchange is None:
            self.arena.update()
        if self.tp is not None:
            self.tp.end_step()
        if self.trace_buf is not None:
            if self.exchange is not None:
                self.exchange.wait_all()
            self._mark(2 + 2 * n)
        ops.check(ops.lib().cxn_counter_add(self.step_counter.data_ptr(), 1, ops.stream()), "counter_add")

    def update_all(self, epoch: int) -> None:
        if self.backend == "ref":
            for us in self.updaters:
                for u in us:
                    u.update(epoch)
        else:
            self.arena.fused_on = False
            self.arena.refresh_hyper(epoch)
            self.arena.update()

    def fuse_update_ok(self) -> bool:
        """whether a step that updates may leave the fully connected weight matrices to their weight-gradient kernel"""
        return bool(self.fuse_update and self.exchange is None and self.tp is None and self.world_size == 1)

    def start_round(self, r: int) -> None:
        for us in self.updaters:
            for u in us:
                u.start_round(r)
        if self.arena is not None:
            for s in self.arena.segments:
                s.param.round = r

    # ------------------------------------------------------------------ access
    def copy_node_data(self, node_id: int) -> torch.Tensor:
        """fp32 (b,c,h,w) copy of a node on the device."""
        n = self.nodes[node_id]
        if self.backend == "ref":
            return n.data.reshape(n.shape).clone()
        return self._node_load(n)

    def _find_weight(self, layer_index: int, tag: str):
        layer = self.connections[layer_index].layer
        for t, w, g in layer.weights():
            if t == tag:
                return layer, w
        return layer, None

    def get_weight(self, layer_index: int, tag: str):
        """(2-D flattened weight, original shape) - reference ``visitor.h:25-91`` GetWeightVisitor."""
        self.sync_weights_from_device()
        _, w = self._find_weight(layer_index, tag)
        if w is None:
            return None, []
        shape = list(w.shape)
        return w.reshape(shape[0], -1).detach().cpu().clone(), shape

    def get_grad(self, layer_index: int, tag: str) -> Optional[torch.Tensor]:
        """accumulated weight gradient in the checkpoint layout (testing / debugging aid)."""
        c = self.connections[layer_index]
        for t, w, g in c.layer.weights():
            if t != tag:
                continue
            if self.arena is not None and c.native:
                if (layer_index, tag) not in self._seg_of:      # a share[tag] connection: the weights are the primary's
                    layer_index = next(j for j, k in enumerate(self.connections) if k.layer is c.layer)
                s = self._seg_of[(layer_index, tag)]
                return c.layer.weight_from_internal(tag, self.arena.g_view(s)).reshape(w.shape).detach().clone()
            return g.detach().clone()
        return None

    def set_weight(self, layer_index: int, weight: torch.Tensor, tag: str) -> None:
        self.sync_weights_from_device()
        _, w = self._find_weight(layer_index, tag)
        if w is None:
            return
        if weight.numel() != w.numel():
            raise ValueError("SetWeight: weight size mismatch")
        w.copy_(weight.reshape(w.shape).to(w.device))
        self.sync_weights_to_device()
