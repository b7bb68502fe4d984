"""Asynchronous parameter server: ``param_server = dist`` with ``update_on_server = 1``.

The reference's multi-node mode (``bin/cxxnet.ps``, ps-lite; reference ``src/nnet/nnet_ps_server.cpp:20-149``,
``src/updater/async_updater-inl.hpp:109-113,200-205``, ``example/MNIST/mpi.conf``): server processes own the weights and a
CPU updater per key (``key = layer * 4 + tag``, ``updater.h:150-173``); a worker pushes a layer's gradient after its
backward pass, the server applies it THE MOMENT IT ARRIVES (no waiting for the other workers - stale gradients are part
of the method) and the worker pulls the new weights before the layer's next forward pass.

Here the roles are ranks of one ``torch.distributed`` job: the last ``num_servers`` ranks are servers, the others train
(the batch is divided among the workers only).  Traffic is point-to-point over a gloo group on host tensors - a
parameter server's wire is the network, so a GPU worker stages gradients and weights through pinned memory, exactly as
the reference's workers copy through the host.  Keys are spread over the servers round robin.

Wire format, worker -> server: a header ``int64[4] = (op, key, numel, flag)`` then, for INIT / PUSH, ``float32[numel]``.
Server -> worker, for PUSH and PULL: ``float32[numel]``, the weights after the update.  Messages between one worker and
one server are ordered; the server takes headers from ANY worker (``dist.recv`` without a source).

==========  =====================================================================================================
``INIT``    first writer wins: the weights of ``key`` if the server has none yet (the reference's ``InitModel``,
            ``nnet_ps_server.cpp:59-82``: "already inited" is not an error); flag = 1 overwrites (a loaded model).
            Acknowledged with the header, so that the sender knows the key exists before it lets other workers ask
``PUSH``    gradient of ``key``: ``updater.update(epoch)`` at once (``nnet_ps_server.cpp:83-89``), reply = new weights
``PULL``    reply = current weights
``DONE``    this worker has finished; a server returns from ``serve()`` when every worker has said so
==========  =====================================================================================================

The learning-rate schedules count UPDATES per key; with W workers a server sees W pushes per global step, so the epoch
it hands the updater is ``pushes // W`` - the schedules advance at the pace of synchronous training.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Tuple

import torch

OP_INIT, OP_PUSH, OP_PULL, OP_DONE = 1, 2, 3, 4
_TAG_ID = {"wmat": 0, "bias": 1}


def encode_key(layer_index: int, tag: str) -> int:
    """reference ``updater.h:150-173`` (EncodeDataKey: layer * kDataKeyStep + tag index)"""
    return layer_index * 4 + _TAG_ID.get(tag, 2)


class Roles:
    """who is what in a job of ``world`` ranks with ``num_servers`` servers (the last ranks)"""

    def __init__(self, world: int, rank: int, num_servers: int):
        if num_servers < 1 or num_servers >= world:
            raise ValueError("update_on_server = 1 needs at least one worker and one server: %d ranks, num_servers = %d"
                             % (world, num_servers))
        self.world, self.rank, self.num_servers = world, rank, num_servers
        self.num_workers = world - num_servers
        self.is_server = rank >= self.num_workers
        self.servers = list(range(self.num_workers, world))

    def server_of(self, key: int) -> int:
        return self.servers[key % self.num_servers]


class ParamServer:
    """server role: one entry per key it owns - flat fp32 weights on the host and a CPU updater built from the
    same configuration the workers read (the reference's UpdaterEntry)"""

    def __init__(self, roles: Roles, group, make_updater):
        """``make_updater(key, w, dw) -> Updater`` builds the updater of a key on the given flat tensors"""
        self.roles, self.group, self.make_updater = roles, group, make_updater
        self.entries: Dict[int, Tuple[torch.Tensor, torch.Tensor, object]] = {}
        self.pushes: Dict[int, int] = {}
        self.stats = {"init": 0, "push": 0, "pull": 0}

    def _entry(self, key: int, numel: int):
        e = self.entries.get(key)
        if e is None:
            w = torch.zeros(numel, dtype=torch.float32)
            dw = torch.zeros(numel, dtype=torch.float32)
            e = self.entries[key] = (w, dw, self.make_updater(key, w, dw))
            self.pushes[key] = 0
        if e[0].numel() != numel:
            raise RuntimeError("parameter server: key %d arrived with %d elements, the server holds %d" % (key, numel, e[0].numel()))
        return e

    def serve(self) -> Dict[str, int]:
        """handle requests until every worker has sent DONE; returns the request counts"""
        import torch.distributed as dist
        done = 0
        header = torch.zeros(4, dtype=torch.int64)
        while done < self.roles.num_workers:
            src = dist.recv(header, group=self.group)
            op, key, numel = int(header[0]), int(header[1]), int(header[2])
            if op == OP_DONE:
                done += 1
                continue
            if op == OP_INIT:
                fresh = key not in self.entries
                w, dw, upd = self._entry(key, numel)
                buf = torch.empty(numel, dtype=torch.float32)
                dist.recv(buf, src=src, group=self.group)
                if fresh or int(header[3]) == 1:        # first writer wins, unless the worker insists (a loaded model)
                    w.copy_(buf)
                    self.stats["init"] += 1
                # acknowledged: the sender may tell the other workers that the key exists only when it does (their
                # requests travel on other connections, which this loop does not order against this one)
                dist.send(header, dst=src, group=self.group)
            elif key not in self.entries:
                raise RuntimeError("parameter server %d: request %d from rank %d for key %d, which nobody has initialised"
                                   % (self.roles.rank, op, src, key))
            elif op == OP_PUSH:
                w, dw, upd = self._entry(key, numel)
                dist.recv(dw, src=src, group=self.group)
                upd.update(self.pushes[key] // self.roles.num_workers)      # (the updater zeroes dw)
                self.pushes[key] += 1
                self.stats["push"] += 1
                dist.send(w, dst=src, group=self.group)
            elif op == OP_PULL:
                w, dw, upd = self._entry(key, numel)
                self.stats["pull"] += 1
                dist.send(w, dst=src, group=self.group)
            else:
                raise RuntimeError("parameter server: unknown request %d from rank %d" % (op, src))
        return dict(self.stats)


class ServerClient:
    """worker side of the protocol; every tensor is a flat fp32 host tensor"""

    def __init__(self, roles: Roles, group):
        self.roles, self.group = roles, group
        self._hdr = torch.zeros(4, dtype=torch.int64)

    def _send_header(self, op: int, key: int, numel: int, dst: int, flag: int = 0) -> None:
        import torch.distributed as dist
        self._hdr[0], self._hdr[1], self._hdr[2], self._hdr[3] = op, key, numel, flag
        dist.send(self._hdr, dst=dst, group=self.group)

    def init(self, key: int, w: torch.Tensor, force: bool = False) -> None:
        import torch.distributed as dist
        dst = self.roles.server_of(key)
        self._send_header(OP_INIT, key, w.numel(), dst, 1 if force else 0)
        dist.send(w, dst=dst, group=self.group)
        ack = torch.zeros(4, dtype=torch.int64)
        dist.recv(ack, src=dst, group=self.group)        # the server has the key now

    def push_pull(self, key: int, grad: torch.Tensor, out_w: torch.Tensor) -> None:
        """gradient up, weights after the server's update down"""
        import torch.distributed as dist
        dst = self.roles.server_of(key)
        self._send_header(OP_PUSH, key, grad.numel(), dst)
        dist.send(grad, dst=dst, group=self.group)
        dist.recv(out_w, src=dst, group=self.group)

    def pull(self, key: int, out_w: torch.Tensor) -> None:
        import torch.distributed as dist
        dst = self.roles.server_of(key)
        self._send_header(OP_PULL, key, out_w.numel(), dst)
        dist.recv(out_w, src=dst, group=self.group)

    def done(self) -> None:
        for s in self.roles.servers:
            self._send_header(OP_DONE, 0, 0, s)


class _Param:
    """one key of a worker: where its gradient is read and its weights are written (arena segment of the b200
    backend, or the (w, dw) pair of a reference-backend updater), plus a flat host staging buffer"""

    def __init__(self, key, n, host, read_grad, write_weight, read_weight):
        self.key, self.n, self.host = key, n, host
        self.read_grad, self.write_weight, self.read_weight = read_grad, write_weight, read_weight


class ServerExchange:
    """Gradient exchange of a WORKER through the parameter servers; the interface the executor and the trainer use
    for every exchange (``exchange.GradExchange``, ``dist_exchange.CollectiveExchange``).

    When a layer's backward pass has finished, its gradients go (b200 backend: device -> pinned host ->) to the server
    that owns the key and the server's weights after ITS update come back into the fp32 masters (the bf16 copies are
    refreshed before the next forward pass - ``UpdateWait``, ``async_updater-inl.hpp:140-143``).  The conversation with
    the server runs on the host in program order: a parameter-server step is bounded by the network, not by launch
    overhead.  Works for both backends: the reference backend's updaters keep their (w, dw) tensors, only the update
    itself happens on the server.
    """

    def __init__(self, net, roles: Roles, group, worker_group=None):
        self.net, self.roles = net, roles
        self.client = ServerClient(roles, group)
        self.wgroup = worker_group
        self.rank, self.world = net.rank, roles.num_workers
        self.device = net.device
        self.shards: Dict = {}
        self.arena = getattr(net, "arena", None) if net.backend == "b200" else None
        cuda = self.device.type == "cuda"
        self.by_layer: Dict[int, List[_Param]] = {}
        self.params: List[_Param] = []

        def host(n):
            t = torch.zeros(n, dtype=torch.float32)
            return t.pin_memory() if cuda else t
        if self.arena is not None:
            a = self.arena
            for s in a.segments:
                h = host(s.n)

                def read_grad(s=s, h=h):
                    h.copy_(a.g_view(s))          # (synchronous device -> host copy on the current stream)
                    if not s.nozero:
                        a.g_view(s).zero_()       # accumulated with atomics: the next step starts from zero
                p = _Param(encode_key(s.layer_index, s.tag), s.n, h, read_grad,
                           lambda s=s, h=h: a.w_view(s).copy_(h, non_blocking=cuda),
                           lambda s=s, h=h: h.copy_(a.w_view(s)))
                self.params.append(p)
                self.by_layer.setdefault(s.layer_index, []).append(p)
        else:
            for li, us in enumerate(net.updaters):
                for u in us:
                    h = host(u.w.numel())

                    def read_grad(u=u, h=h):
                        h.copy_(u.dw.reshape(-1))
                        u.dw.zero_()
                    p = _Param(encode_key(li, u.param.tag), u.w.numel(), h, read_grad,
                               lambda u=u, h=h: u.w.copy_(h.view_as(u.w)),
                               lambda u=u, h=h: h.copy_(u.w.reshape(-1)))
                    self.params.append(p)
                    self.by_layer.setdefault(li, []).append(p)
        self._dirty = False

    def _refresh(self) -> None:
        if self.arena is not None:
            self.arena.refresh_shadow()

    # weights: worker `root` offers its values (first writer wins unless `force`), then everybody takes what the
    # servers hold
    def broadcast_weights(self, root: int = 0, force: bool = False) -> None:
        import torch.distributed as dist
        if self.rank == root:
            for p in self.params:
                p.read_weight()
                self.client.init(p.key, p.host, force)
        # a server handles one worker's messages in order, but not the workers relative to each other: nobody may ask
        # for a key before the root's INIT for it has gone out
        if self.roles.num_workers > 1:
            dist.barrier(group=self.wgroup)
        self.pull_all()

    def pull_all(self) -> None:
        for p in self.params:
            self.client.pull(p.key, p.host)
            p.write_weight()
        self._refresh()
        if self.device.type == "cuda":
            torch.cuda.synchronize(self.device)

    def begin_step(self) -> None:
        pass

    def layer_done(self, layer_index: int) -> None:
        for p in self.by_layer.get(layer_index, ()):
            p.read_grad()
            self.client.push_pull(p.key, p.host, p.host)
            p.write_weight()
            self._dirty = True

    def wait_layer(self, layer_index: int) -> None:
        pass            # the copies above are stream ordered before anything the next forward pass launches

    def wait_all(self) -> None:
        if self._dirty:
            self._refresh()
            self._dirty = False

    def publish_pending(self) -> None:
        self.wait_all()

    def after_replayed_step(self) -> None:
        pass

    def gather_master(self) -> None:
        self.wait_all()

    def fc_backward(self, layer, need_update: bool) -> None:      # pragma: no cover - never planned on this path
        raise RuntimeError("fullc_gather shards are not used with the parameter server")

    def check_replicas(self) -> float:
        """largest |w - server's w| (asynchronous workers are not replicas of each other; right after a pull with no
        other worker pushing this is 0)"""
        self.wait_all()
        worst = 0.0
        for p in self.params:
            p.read_weight()
            mine = p.host.clone()
            self.client.pull(p.key, p.host)
            worst = max(worst, float((mine - p.host).abs().max()))
        return worst

    def finish(self) -> None:
        self.client.done()
