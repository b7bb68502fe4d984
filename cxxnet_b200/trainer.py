"""Net trainer: the ``INetTrainer`` surface (reference ``src/nnet/nnet.h:18-92``), implemented in the
one-process-per-GPU model.

Parity with ``CXXNetThreadTrainer`` (reference ``src/nnet/nnet_impl-inl.hpp:16-455``):

* ``dev`` parser (``cpu``, ``gpu``, ``gpu:0,1,2``, ``gpu:0-3``), ``batch_size`` is the GLOBAL
  batch split ``ceil(batch/ndev)`` per device (``:147``), surplus devices dropped (``:344-354``);
* ``update_period`` gradient accumulation, ``eval_train`` train metrics, ``metric`` /
  ``metric[label,node]`` sets, ``seed``, ``param_server``;
* ``update`` / ``predict`` / ``extract_feature`` / ``evaluate`` / ``get_weight`` /
  ``set_weight`` / ``save_model`` / ``load_model`` / ``copy_model_from``.

Where the reference runs one pthread per GPU inside one process and exchanges gradients
through mshadow-ps, here each rank of a ``torch.distributed`` job owns one device and its
slice ``[rank*step, (rank+1)*step)`` of every global batch; gradients are exchanged by the
fused NVLink kernel (``parallel/exchange.py``).  A single process given several devices
re-launches itself with one worker per device (``parallel/launch.py``).
"""
from __future__ import annotations

import io
import os
import re
import sys
from typing import BinaryIO, List, Optional, Tuple

import numpy as np
import torch

from . import model_io as mio
from . import ops
from .config import ConfigPairs
from .executor import NeuralNet
from .io.data import DataBatch, IIterator
from .metrics import MetricSet
from .netconfig import NetConfig


def parse_devices(val: str) -> Tuple[str, List[int]]:
    """``cpu`` | ``gpu`` | ``gpu:0,1,2`` | ``gpu:0-3`` (reference ``nnet_impl-inl.hpp:32-51``)."""
    kind = "gpu" if val.startswith("gpu") else "cpu"
    devs: List[int] = []
    if ":" in val:
        spec = val.split(":", 1)[1]
        m = re.match(r"^(\d+)-(\d+)$", spec)
        if m:
            devs = list(range(int(m.group(1)), int(m.group(2)) + 1))
        else:
            for p in spec.split(","):
                if p:
                    devs.append(int(p))
    return kind, devs


class NetTrainer:
    def __init__(self, net_type: int = 0):
        self.net_type = net_type
        self.cfg: ConfigPairs = []
        self.net_cfg = NetConfig()
        self.batch_size = 100
        self.update_period = 1
        self.sample_counter = 0
        self.eval_train = 1
        self.epoch_counter = 0
        self.seed = 0
        self.silent = 0
        self.type_pserver = "UNSPECIFIED"
        self.dev_kind = "cpu"
        self.devices: List[int] = []
        self.backend_override: Optional[str] = None
        self.metric = MetricSet()
        self.train_metric = MetricSet()
        self.eval_nodes: List[Tuple[str, int]] = []
        self.eval_req: List[int] = []
        self.net: Optional[NeuralNet] = None
        self.rank, self.world_size = 0, 1
        self.step = 0
        self._stage = None
        # cuda_graph = 1: after a few eager steps one training step (forward, metric, backward, gradient
        # exchange, update) is captured into a CUDA graph per input staging slot and replayed; the host
        # then only stages the batch and launches one graph per step
        self.cuda_graph = 0
        self.test_on_server = 0
        self.update_on_server = 0
        self.num_servers = 1
        self.is_server = False
        self.ps_roles = self.ps_group = self.ps_wgroup = self.ps_server = None
        self.exchange_kind = "nvlink"
        self.check_nan = 0
        self.layer_timing = 0
        self.step_trace = int(os.environ.get("CXN_STEP_TRACE", "0"))
        self.graph_warmup = 3
        self._graphs = {}            # staging slot -> (torch.cuda.CUDAGraph, launches inside)
        self._graph_rows = -1
        self._graph_unsafe = None
        self._graph_steps = 0
        self._graph_labels: Optional[List[torch.Tensor]] = None

    # ------------------------------------------------------------------ configuration
    def set_param(self, name: str, val: str) -> None:
        if name == "dev":
            self.dev_kind, self.devices = parse_devices(val)
        if name == "batch_size": self.batch_size = int(val)
        if name == "update_period": self.update_period = int(val)
        if name == "eval_train": self.eval_train = int(val)
        if name == "seed": self.seed = int(val)
        if name == "silent": self.silent = int(val)
        if name == "param_server": self.type_pserver = val
        # param_server = dist with update_on_server = 1: asynchronous parameter server (parallel/ps_server.py) - the
        # last `num_servers` ranks own the weights and the updaters, the others train (reference
        # async_updater-inl.hpp:109-113, nnet_ps_server.cpp)
        if name == "update_on_server": self.update_on_server = int(val)
        if name == "num_servers": self.num_servers = int(val)
        if name == "backend": self.backend_override = val
        if name == "cuda_graph": self.cuda_graph = int(val)
        # reference async_updater-inl.hpp:148-153: replicate the weights on the server and assert equality every
        # round.  Here: every round all ranks compare their replicas with rank 0's (collective).
        if name == "test_on_server": self.test_on_server = int(val)
        if name == "layer_timing": self.layer_timing = int(val)     # per-layer timers, see utils/timing.py
        if name == "step_trace": self.step_trace = int(val)         # device timestamps between layers (executor.enable_trace)
        # check_nan = 1: every round starts with a look at the weights; a NaN / Inf stops the run and names the
        # layer (the reference only notices inside the logloss metric, src/utils/metric.h:128)
        if name == "check_nan": self.check_nan = int(val)
        # Parameter-server tuning keys of the reference (async_updater-inl.hpp:155-184, fullc_layer-inl.hpp) that
        # have nothing to act on here - there is no server, pulls are not deferred, arrays are not chunked by
        # the transport: accepted and ignored, like every key a consumer does not know.
        #   pull_at_backprop, bigarray_bound, init_on_worker, force_contiguous
        if name.startswith("metric"):
            m = re.match(r"metric\[([^,]+),([^\]]+)\]", name)
            if m:
                self.metric.add_metric(val, m.group(1))
                self.train_metric.add_metric(val, m.group(1))
                self.eval_nodes.append((m.group(2), 0))
            else:
                # anything else, including ``metric[label]``, evaluates field "label" on the
                # last node, as the reference's sscanf does (nnet_impl-inl.hpp:57-67)
                self.metric.add_metric(val, "label")
                self.train_metric.add_metric(val, "label")
                self.eval_nodes.append(("", -1))
        self.cfg.append((name, val))

    # ------------------------------------------------------------------ device / process layout
    def _resolve_devices(self) -> None:
        import torch.distributed as dist
        self.exchange_kind = "nvlink"
        if self.type_pserver == "dist":
            # The reference's multi-node mode (ps-lite servers, example/MNIST/mpi.conf).  Here: the ranks of a
            # torch.distributed job launched over several nodes, exchanging through collectives
            # (parallel/dist_exchange.py) instead of peer-mapped memory.
            if not (dist.is_available() and dist.is_initialized()):
                raise ValueError("param_server=dist needs a torch.distributed job: launch one rank per device "
                                 "with torchrun (--nnodes N --nproc-per-node G) and run `python -m cxxnet_b200 conf`")
            self.exchange_kind = "collective"
        if dist.is_available() and dist.is_initialized() and self.type_pserver != "NONE":
            # param_server = NONE keeps this trainer private to its process (no gradient exchange),
            # as in the reference when a single device is used (nnet_impl-inl.hpp:378-381)
            self.world_size, self.rank = dist.get_world_size(), dist.get_rank()
        if self.type_pserver == "dist" and self.update_on_server:
            from .parallel.ps_server import Roles
            if self.ps_roles is None:
                self.ps_roles = Roles(dist.get_world_size(), dist.get_rank(), self.num_servers)
                # (group creation is collective: every rank, once, in this order.)  Requests and weights travel point
                # to point on host tensors, whatever the job's default backend is; the workers keep a group of their
                # own for the metric / prediction collectives - the servers are busy serving.
                self.ps_group = dist.new_group(backend="gloo")
                self.ps_wgroup = dist.new_group(ranks=list(range(self.ps_roles.num_workers)), backend="gloo")
            self.exchange_kind = "server"
            self.world_size = self.ps_roles.num_workers          # the batch is divided among the workers
            self.is_server = self.ps_roles.is_server
            if self.is_server:
                self.dev_kind = "cpu"                            # a server's updaters run on the host, as the reference's
        if self.world_size > 1 and self.dev_kind == "gpu" and self.exchange_kind == "nvlink":
            from .parallel.dist_exchange import ranks_share_one_host
            # peer-mapped memory exists inside one machine only; CXN_EXCHANGE=collective selects the library
            # path on one node too (the arm the fused kernels are compared against)
            if os.environ.get("CXN_EXCHANGE", "") == "collective" or not ranks_share_one_host():
                self.exchange_kind = "collective"
        ndev = max(len(self.devices), 1)
        if self.world_size > 1:
            ndev = self.world_size
        step = max((self.batch_size + ndev - 1) // ndev, 1)
        used = ndev
        while step * (used - 1) >= self.batch_size and used > 1:
            used -= 1
        if used < ndev and self.silent == 0 and self.rank == 0:
            print("Warning: The number of devices is induce mini-batch=%u\n"
                  "We can equally use %u devices to cover the batch_size" % (step, used))
        self.step = step
        self.active_ranks = used
        if self.dev_kind == "gpu":
            if not torch.cuda.is_available():
                raise RuntimeError("dev=gpu requested but no CUDA device is available")
            if self.world_size > 1:
                local = int(os.environ.get("LOCAL_RANK", self.rank))
                idx = self.devices[local] if local < len(self.devices) else local
            elif self.type_pserver == "NONE" and len(self.devices) <= 1:
                idx = self.devices[0] if self.devices else torch.cuda.current_device()
            else:
                idx = self.devices[0] if self.devices else torch.cuda.current_device()
                if len(self.devices) > 1:
                    raise RuntimeError(
                        "dev lists %d GPUs but this is a single process: launch with "
                        "`python -m cxxnet_b200.parallel.launch` / torchrun (one rank per GPU)" % len(self.devices))
            torch.cuda.set_device(idx)
            self.device = torch.device("cuda", idx)
            self.backend = self.backend_override or "b200"
        else:
            self.device = torch.device("cpu")
            self.backend = "ref"
        if self.backend not in ("ref", "b200"):
            raise ValueError("unknown backend %s" % self.backend)

    def _init_net(self) -> None:
        assert self.net is None, "net must be empty before this"
        self.net_cfg.configure(self.cfg)
        self._resolve_devices()
        if self.is_server:
            self._init_server()
            return
        self.net = NeuralNet(self.net_cfg, self.step, self.rank + self.seed * 100, self.device, self.backend,
                             silent=self.silent or self.rank != 0, world_size=self.world_size, rank=self.rank,
                             exchange_kind=self.exchange_kind)
        if self.exchange_kind == "server" or (self.world_size > 1 and self.exchange_kind == "collective"):
            self.cuda_graph = 0          # library collectives / server round trips stay outside captured steps
        if self.layer_timing:
            from .utils.timing import LayerTimer
            def names():
                return ["%s:%s" % (c.layer.type_name, self.net_cfg.layers[i].name or i)
                        for i, c in enumerate(self.net.connections)]
            self.net.timer = LayerTimer(names, self.device.type == "cuda")
            self.cuda_graph = 0          # a captured step cannot be bracketed
        if self.silent == 0 and self.rank == 0:
            print("finish initialization with %d devices" % max(self.world_size, 1))
        self.eval_req = []
        for name, flag in self.eval_nodes:
            if flag < 0:
                self.eval_req.append(self.net_cfg.num_nodes - 1)
            else:
                if name not in self.net_cfg.node_name_map:
                    raise ValueError("Cannot find node name: %s\n" % name)
                self.eval_req.append(self.net_cfg.node_name_map[name])

    def _finish_init(self) -> None:
        if self.is_server:
            return
        self.net.init_updaters()
        self.net.init_nodes()
        if self.step_trace and self.net.backend == "b200":
            self.net.enable_trace()
        if self.exchange_kind == "server":
            from .parallel.ps_server import ServerExchange
            ex = ServerExchange(self.net, self.ps_roles, self.ps_group, self.ps_wgroup)
            self.net.exchange = ex            # (both backends: the executor reports finished layers to it)

    # ------------------------------------------------------------------ parameter-server role
    def _init_server(self) -> None:
        """this rank is a server: no net, one CPU updater per key built from the configuration the workers read
        (reference nnet_ps_server.cpp:59-82 InitModel)"""
        from .parallel.ps_server import ParamServer
        from .updaters import create_updater
        cfg = self.net_cfg
        tags = {0: "wmat", 1: "bias"}

        def make_updater(key, w, dw):
            li, tag = key // 4, tags.get(key % 4, "wmat")
            if li >= len(cfg.layers):
                raise RuntimeError("parameter server: key %d names layer %d of a net with %d layers" % (key, li, len(cfg.layers)))
            u = create_updater(cfg.updater_type, w, dw, tag)
            for k, v in list(cfg.defcfg) + list(cfg.layercfg[li]):
                u.set_param(k, v)
            u.param.silent = 1
            u.init()
            return u
        self.ps_server = ParamServer(self.ps_roles, self.ps_group, make_updater)

    def serve(self):
        """server ranks: answer the workers until all of them have finished; returns the request counts"""
        assert self.is_server and self.ps_server is not None, "serve() is for the server ranks of update_on_server = 1"
        return self.ps_server.serve()

    def finish(self) -> None:
        """end of the job: a worker of a parameter-server run tells the servers it is done"""
        if self.exchange_kind == "server" and not self.is_server and self.net is not None and self.net.exchange is not None:
            self.net.exchange.finish()

    # ------------------------------------------------------------------ model
    def init_model(self) -> None:
        self._init_net()
        if self.is_server:
            return
        if self.world_size > 1:
            # every rank draws the same initial weights (the reference serialises device 0's model
            # to the others, nnet_impl-inl.hpp:70-81); the exchange re-broadcasts rank 0's arena.
            self.net.rng.manual_seed(self.seed * 100)
        if self.is_server:
            return
        self.net.init_model()
        self._finish_init()
        if self.net.exchange is not None:
            self.net.exchange.broadcast_weights()

    def save_model(self, fo: BinaryIO) -> None:
        self.net_cfg.save_net(fo)
        mio.write_i64(fo, self.epoch_counter)
        blob = io.BytesIO()
        self.net.save_model(blob)
        mio.write_str(fo, blob.getvalue())

    def load_model(self, fi: BinaryIO) -> None:
        self.net_cfg.load_net(fi)
        self.epoch_counter = mio.read_i64(fi)
        self.net = None
        self._init_net()
        blob = mio.read_bytes(fi)
        if self.is_server:
            return
        self.net.load_model(io.BytesIO(blob))
        self._finish_init()
        if self.exchange_kind == "server":
            self.net.exchange.broadcast_weights(force=True)      # the servers take the loaded weights

    def copy_model_from(self, fi: BinaryIO) -> None:
        """finetune: copy layers whose names match (reference ``nnet_impl-inl.hpp:101-134``)."""
        self.net = None
        self.init_model()
        if self.is_server:
            return
        old_cfg = NetConfig()
        old_cfg.load_net(fi)
        mio.read_i64(fi)
        self.epoch_counter = 0
        old_net = NeuralNet(old_cfg, 1, 0, torch.device("cpu"), "ref", silent=1)
        old_net.load_model(io.BytesIO(mio.read_bytes(fi)), infer_shapes=False)
        for i, ol in enumerate(old_cfg.layers):
            for j, nl in enumerate(self.net_cfg.layers):
                if ol.name == nl.name and ol.name != "":
                    print("Copying layer %s" % ol.name)
                    buf = io.BytesIO()
                    old_net.connections[i].layer.save_model(buf)
                    buf.seek(0)
                    self.net.copy_layer(j, buf)
        if self.exchange_kind == "server":
            self.net.exchange.broadcast_weights(force=True)

    def find_non_finite(self):
        """(layer name, tag) of the first weight tensor holding a NaN or an infinity, else None"""
        net = self.net
        if net.backend == "b200" and net.arena is not None:
            if net.exchange is not None:
                net.exchange.wait_all()
            if bool(torch.isfinite(net.arena.w[:max(net.arena.total, 1)]).all()):
                return None
            for s in net.arena.segments:
                if not bool(torch.isfinite(net.arena.w_view(s)).all()):
                    return self.net_cfg.layers[s.layer_index].name or str(s.layer_index), s.tag
            return None
        for i, c in enumerate(net.connections):
            if c.type == 0:          # shared layer: its weights belong to the primary
                continue
            for tag, w, _ in c.layer.weights():
                if not bool(torch.isfinite(w).all()):
                    return self.net_cfg.layers[i].name or str(i), tag
        return None

    def start_round(self, r: int) -> None:
        self.net.start_round(r)
        if self.check_nan:
            bad = self.find_non_finite()
            if bad is not None:
                raise RuntimeError("check_nan: non-finite weights in layer %s (%s) at the start of round %d - lower the "
                                   "learning rate or set clip_gradient" % (bad[0], bad[1], r))
        ex = getattr(self.net, "exchange", None)
        if self.test_on_server and ex is not None:
            d = ex.check_replicas()
            if d != 0.0:
                raise RuntimeError("test_on_server: weight replicas differ from rank 0 by up to %g at round %d" % (d, r))

    def timing_report(self, reset: bool = True) -> str:
        """per-layer forward / backward time since the last report (`layer_timing = 1`), else an empty string"""
        tm = self.net.timer if self.net is not None else None
        if tm is None:
            return ""
        text = tm.report()
        if reset:
            tm.reset()
        return text

    # ------------------------------------------------------------------ batch slicing
    def _slice(self, n: int) -> Tuple[int, int]:
        begin = min(self.rank * self.step, n)
        end = min((self.rank + 1) * self.step, n)
        return begin, end

    def _label_fields(self, label: torch.Tensor, begin: int, end: int, device) -> List[torch.Tensor]:
        out = []
        for a, b in self.net_cfg.label_range:
            out.append(label[begin:end, a:b].to(device, non_blocking=True).float().contiguous())
        return out

    def _as_tensor(self, x) -> torch.Tensor:
        if isinstance(x, torch.Tensor):
            return x
        return torch.from_numpy(np.ascontiguousarray(x, dtype=np.float32))

    # ------------------------------------------------------------------ training
    def update(self, batch: DataBatch) -> None:
        """one mini-batch: forward, backprop, (every ``update_period`` calls) weight update
        (reference ``nnet_impl-inl.hpp:141-185``)."""
        net = self.net
        if batch.batch_size > self.batch_size:
            raise ValueError("update: the batch holds %d instances but the net was built for batch_size = %d"
                             % (batch.batch_size, self.batch_size))
        if self.cuda_graph and self._graph_applicable(batch):
            self._update_graph(batch)
            return
        need_update = (self.sample_counter + 1) % self.update_period == 0
        data = self._as_tensor(batch.data)
        label = self._as_tensor(batch.label)
        begin, end = self._slice(batch.batch_size)
        net.label_info.fields = self._label_fields(label, begin, end, self.device)
        net.accumulate_grads = self.sample_counter != 0
        net.input_norm = getattr(batch, "norm", None)
        net.set_input(data[begin:end], [self._as_tensor(e)[begin:end] for e in batch.extra_data])
        net.forward(True)
        self._eval_train()
        net.backprop(False, need_update, self.epoch_counter)
        self.sample_counter += 1
        if self.sample_counter >= self.update_period:
            self.sample_counter = 0
            self.epoch_counter += 1

    def _eval_train(self) -> None:
        """running train metric on the forward pass just done.  `error` on a native loss node is one kernel
        that accumulates into the metric's device scalar; anything else goes through the torch reductions."""
        if self.eval_train == 0 or not self.eval_req:
            return
        net = self.net
        fields, fmap = net.label_info.fields, self.net_cfg.label_name_map
        ms = self.train_metric
        fast = net.backend == "b200" and all(
            type(e).__name__ == "MetricError" and f in fmap and net.nodes[nid].loss_input and net.nodes[nid].f32 is not None
            and fields[fmap[f]].is_cuda and fields[fmap[f]].dtype == torch.float32 and fields[fmap[f]].is_contiguous()
            for e, f, nid in zip(ms.evals, ms.label_fields, self.eval_req))
        if not fast:
            scores = [self._scores2d(net.copy_node_data(i)) for i in self.eval_req]
            ms.add_eval(scores, fields, fmap)
            return
        for e, f, nid in zip(ms.evals, ms.label_fields, self.eval_req):
            n = net.nodes[nid]
            rows, cols = n.shape[0], n.shape[3]
            lab = fields[fmap[f]]
            if e.sum_metric is None or e.sum_metric.device != n.f32.device:
                e.sum_metric = torch.zeros((), dtype=torch.float64, device=n.f32.device)
            ops.check(ops.lib().cxn_metric_error(n.f32.data_ptr(), n.f32.shape[1], lab.data_ptr(), lab.shape[1], rows, cols,
                                                 e.sum_metric.data_ptr(), ops.stream()), "metric_error")
            e.cnt_inst += rows

    # ------------------------------------------------------------------ captured training step
    def _graph_applicable(self, batch: DataBatch) -> bool:
        net = self.net
        if net is None or net.backend != "b200" or self.update_period != 1 or batch.extra_data:
            return False
        if self._graph_unsafe is None:     # layers whose kernel arguments change on the host between steps
            self._graph_unsafe = any(getattr(c.layer, "graph_unsafe", False) or not c.native for c in net.connections)
        if self._graph_unsafe:
            return False
        begin, end = self._slice(batch.batch_size)
        # only full-size slices are replayed (a smaller last batch takes the eager path)
        return end - begin == net.max_batch

    def _step_body(self, slot: int) -> None:
        """the part of a training step that lives inside the graph"""
        net = self.net
        net.convert_staged(slot)
        net.forward(True)
        self._eval_train()
        net.backprop(False, True, self.epoch_counter)
        if net.exchange is not None:
            net.exchange.wait_all()          # join the exchange stream: the step is one unit

    def _update_graph(self, batch: DataBatch) -> None:
        net = self.net
        data = self._as_tensor(batch.data)
        label = self._as_tensor(batch.label)
        begin, end = self._slice(batch.batch_size)
        rows = end - begin
        net.adjust_batch_size(rows)
        # persistent label fields (the loss layers and the metric read these addresses)
        if self._graph_labels is None:
            self._graph_labels = [torch.zeros((rows, b - a), dtype=torch.float32, device=self.device)
                                  for a, b in self.net_cfg.label_range]
        for t, (a, b) in zip(self._graph_labels, self.net_cfg.label_range):
            t.copy_(label[begin:end, a:b], non_blocking=True)
        net.label_info.fields = self._graph_labels
        net.accumulate_grads = False
        net.input_norm = getattr(batch, "norm", None)
        slot = net.stage_input(data[begin:end])
        # host half of the per-step hyper parameters (rewrites the pinned table only when it changed)
        net.arena.fused_on = net.fuse_update_ok()
        net.arena.prepare_hyper(self.epoch_counter, slot)
        net.hyper_prepared, net.hyper_slot = True, slot
        try:
            norm = net.input_norm       # baked into a captured conversion kernel's arguments
            gkey = (slot, net.stage_dtype[slot], None if norm is None else (tuple(float(v) for v in norm[0]), float(norm[1])))
            entry = self._graphs.get(gkey)
            if entry is not None:
                g, launches, cnt = entry
                g.replay()
                ops.lib().cxn_launch_count_add(launches)
                for e in self.train_metric.evals:
                    e.cnt_inst += cnt
                if net.exchange is not None:
                    net.exchange.after_replayed_step()
            elif self._graph_steps < self.graph_warmup:
                self._step_body(slot)
            else:
                try:
                    self._capture(slot, rows, gkey)
                except Exception as exc:      # a layer that cannot be captured (host round trips): stay eager
                    import warnings
                    warnings.warn("cuda_graph: capture failed (%r), continuing without graphs" % (exc,))
                    self.cuda_graph = 0
                    self._graphs.clear()
                    torch.cuda.synchronize(self.device)
                    self._step_body(slot)
        finally:
            net.hyper_prepared = False
        net.arena.mark_hyper_consumed(slot)
        net.release_stage(slot)
        self._graph_steps += 1
        self.epoch_counter += 1

    def _capture(self, slot: int, rows: int, gkey) -> None:
        """capture one step for this staging slot; the capture itself does not execute, so the step is
        replayed once right away"""
        net = self.net
        if net.exchange is not None:
            net.exchange.wait_all()
            net.exchange.after_replayed_step()     # the captured forward must contain the deferred publishes
        torch.cuda.synchronize(self.device)
        before = ops.launch_count()
        cnt0 = [e.cnt_inst for e in self.train_metric.evals]
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            self._step_body(slot)
        launches = ops.launch_count() - before
        cnt = (self.train_metric.evals[0].cnt_inst - cnt0[0]) if self.train_metric.evals else 0
        for e, c in zip(self.train_metric.evals, cnt0):
            e.cnt_inst = c
        self._graphs[gkey] = (g, launches, cnt)
        ops.lib().cxn_launch_count_add(-launches & 0xFFFFFFFFFFFFFFFF)   # the captured launches did not run
        g.replay()
        ops.lib().cxn_launch_count_add(launches)
        for e in self.train_metric.evals:
            e.cnt_inst += cnt
        if net.exchange is not None:
            net.exchange.after_replayed_step()

    @staticmethod
    def _scores2d(t: torch.Tensor) -> torch.Tensor:
        return t.reshape(t.shape[0], -1)

    # ------------------------------------------------------------------ inference
    def _forward_to(self, node_ids: List[int], batch: DataBatch) -> List[torch.Tensor]:
        """forward and return the requested nodes as host fp32 (batch_size, c, h, w) tensors
        assembled over all ranks (reference ``nnet_impl-inl.hpp:300-325``)."""
        net = self.net
        if batch.batch_size > self.batch_size:
            # more instances than the net was sized for (an ndarray handed to predict / extract): go through it in
            # pieces of batch_size instead of dropping the rest (the reference stops with an assertion here)
            pieces = []
            for lo in range(0, batch.batch_size, self.batch_size):
                hi = min(lo + self.batch_size, batch.batch_size)
                part = DataBatch()
                part.data = self._as_tensor(batch.data)[lo:hi]
                part.extra_data = [self._as_tensor(e)[lo:hi] for e in batch.extra_data]
                part.norm = getattr(batch, "norm", None)
                part.batch_size = hi - lo
                pieces.append(self._forward_to(node_ids, part))
            return [torch.cat([p[i] for p in pieces], dim=0) for i in range(len(node_ids))]
        data = self._as_tensor(batch.data)
        begin, end = self._slice(batch.batch_size)
        net.input_norm = getattr(batch, "norm", None)
        net.set_input(data[begin:end], [self._as_tensor(e)[begin:end] for e in batch.extra_data])
        net.forward(False)
        outs = []
        for nid in node_ids:
            t = net.copy_node_data(nid)
            if self.world_size > 1:
                t = self._gather_rows(t, batch.batch_size)
            outs.append(t)
        return outs

    def _gather_rows(self, t: torch.Tensor, total: int) -> torch.Tensor:
        import torch.distributed as dist
        full = torch.zeros((self.world_size * self.step,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
        pad = torch.zeros((self.step,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
        pad[: t.shape[0]] = t
        if self.ps_wgroup is not None:       # parameter-server run: the workers' own (gloo) group, through the host
            parts = [torch.zeros_like(pad, device="cpu") for _ in range(self.world_size)]
            dist.all_gather(parts, pad.cpu(), group=self.ps_wgroup)
            return torch.cat(parts)[:total].to(t.device)
        dist.all_gather_into_tensor(full, pad)
        return full[:total]

    def predict(self, batch: DataBatch) -> np.ndarray:
        """argmax (or the raw value when the output width is 1) per instance (reference ``:186-199``)."""
        last = len(self.net.nodes) - 1
        (t,) = self._forward_to([last], batch)
        t = t.reshape(t.shape[0], -1)
        if t.shape[1] != 1:
            mx = t.max(dim=1, keepdim=True).values
            pred = (t == mx).float().argmax(dim=1).float()
        else:
            pred = t[:, 0]
        return pred.cpu().numpy().astype(np.float32)

    def node_id_from_name(self, node_name: str) -> int:
        m = re.match(r"top\[-(\d+)\]", node_name)
        nnode = len(self.net.nodes)
        if m:
            offset = int(m.group(1))
            if not (1 <= offset <= nnode):
                raise ValueError("ExtractFeature: offset must be within num_node range")
            return nnode - offset
        if node_name not in self.net_cfg.node_name_map:
            raise ValueError("ExtractFeature: Cannot find node name: %s" % node_name)
        return self.net_cfg.node_name_map[node_name]

    def extract_feature(self, batch: DataBatch, node_name: str) -> np.ndarray:
        (t,) = self._forward_to([self.node_id_from_name(node_name)], batch)
        return t.cpu().numpy().astype(np.float32)

    def evaluate(self, iter_eval: Optional[IIterator], data_name: str) -> str:
        ret = ""
        if self.eval_train != 0:
            self._reduce_metric(self.train_metric)
            ret += self.train_metric.print("train")
            self.train_metric.clear()
        if iter_eval is None:
            return ret
        self.metric.clear()
        iter_eval.before_first()
        while iter_eval.next():
            batch = iter_eval.value()
            outs = self._forward_to(self.eval_req, batch)
            n = batch.batch_size - batch.num_batch_padd
            label = self._as_tensor(batch.label)
            fields = [label[:n, a:b].float() for a, b in self.net_cfg.label_range]
            scores = [o[:n].reshape(n, -1) for o in outs]
            self.metric.add_eval(scores, fields, self.net_cfg.label_name_map)
        ret += self.metric.print(data_name)
        return ret

    def _reduce_metric(self, ms: MetricSet) -> None:
        if self.world_size <= 1:
            return
        import torch.distributed as dist
        for e in ms.evals:
            s = e.sum_metric if e.sum_metric is not None else torch.zeros((), dtype=torch.float64, device=self.device)
            v = torch.stack([s.to(self.device).double(), torch.tensor(float(e.cnt_inst), dtype=torch.float64,
                                                                       device=self.device)])
            if self.ps_wgroup is not None:
                vc = v.cpu()
                dist.all_reduce(vc, group=self.ps_wgroup)
                v = vc.to(self.device)
            else:
                dist.all_reduce(v)
            if e.sum_metric is not None:
                e.sum_metric.copy_(v[0])          # in place: a captured training step accumulates into this tensor
            else:
                e.sum_metric = v[0].clone()
            e.cnt_inst = int(v[1].item())

    # ------------------------------------------------------------------ weights
    def set_weight(self, weight, layer_name: str, tag: str) -> None:
        if tag not in ("bias", "wmat"):
            raise ValueError("NNet.SetWeight: weight tag can only be bias or wmat")
        idx = self.net_cfg.get_layer_index(layer_name)
        self.net.set_weight(idx, self._as_tensor(weight), tag)

    def get_weight(self, layer_name: str, tag: str):
        if tag not in ("bias", "wmat"):
            raise ValueError("NNet.GetWeight: weight tag can only be bias or wmat")
        idx = self.net_cfg.get_layer_index(layer_name)
        w, shape = self.net.get_weight(idx, tag)
        if w is None:
            return None, []
        return w.numpy(), shape


def create_net(net_type: int = 0) -> NetTrainer:
    """reference ``nnet.h:99-100`` CreateNet."""
    return NetTrainer(net_type)
