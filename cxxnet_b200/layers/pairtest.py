"""``pairtest-<master>-<slave>``: run two implementations of one layer on the same input and
report their relative error (reference ``src/layer/pairtest_layer-inl.hpp:14-200``).

The master runs on the real nodes, the slave on shadow copies; outputs are compared in
forward, weight gradients and input gradients in backward.  Metric: relative L1 error
``sum|m-s| / sum|m|``; a message goes to stderr when it exceeds 1e-5 or is NaN - the run is
never failed.  ``master:`` / ``slave:`` prefixed keys go to one side only.  Results are also kept
in ``self.reports`` so tests can assert on them.

Differences from the reference, on purpose: the reference hands BOTH weight sets to the updater
(``ApplyVisitor`` visits master and slave, ``:120-123``) and lets them evolve independently; here
only the master's weights are trained and the slave's are re-synchronised from them before every
forward, so each step is an independent kernel-vs-kernel comparison.  ``save_model`` writes master
then slave like the reference (``:143-146``); the reference's ``LoadModel`` calls
``slave->SaveModel`` (``:147-150``, a bug) - here both are read back.
"""
from __future__ import annotations

import sys
from typing import List

import torch

from .base import Layer, Node


class PairTestLayer(Layer):
    type_name = "pairtest"
    allow_sharing = False
    tolerance = 1e-5

    def __init__(self, master: Layer, slave: Layer):
        super().__init__()
        self.master = master
        self.slave = slave
        self.snodes_in: List[Node] = []
        self.snodes_out: List[Node] = []
        self.reports: List[tuple] = []

    def set_param(self, name, val):
        if name.startswith("master:"):
            self.master.set_param(name[7:], val)
        elif name.startswith("slave:"):
            self.slave.set_param(name[6:], val)
        else:
            self.master.set_param(name, val)
            self.slave.set_param(name, val)

    def to_device(self, device):
        super().to_device(device)
        self.master.to_device(device)
        self.slave.to_device(device)

    def weights(self):
        return self.master.weights()

    def init_connection(self, nodes_in, nodes_out):
        for l in (self.master, self.slave):
            l.rng, l.label_info, l.layer_index = self.rng, self.label_info, self.layer_index
        self.master.init_connection(nodes_in, nodes_out)
        self.snodes_in = [Node(n.name, n.index) for n in nodes_in]
        for s, n in zip(self.snodes_in, nodes_in):
            s.shape = list(n.shape)
        if self.master.self_loop:
            self.snodes_out = self.snodes_in
        else:
            self.snodes_out = [Node(n.name, n.index) for n in nodes_out]
        self.slave.init_connection(self.snodes_in, self.snodes_out)
        for s, n in zip(self.snodes_out, nodes_out):
            if s.shape != n.shape:
                raise ValueError("PairTestLayer: master and slave disagree on the output shape")

    def _sync(self):
        for (_, mw, _), (tag, sw, sg) in zip(self.master.weights(), self.slave.weights()):
            if sw.device != mw.device:
                self.slave.rebind(tag, sw.to(mw.device), sg.to(mw.device))
                sw = dict((t, w) for t, w, _ in self.slave.weights())[tag]
            sw.copy_(mw.reshape(sw.shape))

    def init_model(self):
        self.master.init_model()
        self.slave.init_model()
        self._sync()

    def save_model(self, fo):
        self._sync()
        self.master.save_model(fo)
        self.slave.save_model(fo)

    def load_model(self, fi):
        self.master.load_model(fi)
        self.slave.load_model(fi)

    def _cmp(self, what: str, m: torch.Tensor, s: torch.Tensor):
        m, s = m.detach().float().reshape(-1), s.detach().float().reshape(-1)
        denom = m.abs().sum().item()
        diff = (m - s).abs().sum().item()
        rel = diff / denom if denom > 0 else diff
        self.reports.append((what, rel))
        if not (rel <= self.tolerance):
            sys.stderr.write("PairTest[%s-%s] %s: relative error %g\n" % (
                self.master.type_name, self.slave.type_name, what, rel))

    def forward(self, is_train, nodes_in, nodes_out):
        self._sync()
        for s, n in zip(self.snodes_in, nodes_in):
            s.shape = list(n.shape)
            s.data = n.data.clone()
        for s, n in zip(self.snodes_out, nodes_out):
            s.shape = list(n.shape)
        self.master.forward(is_train, nodes_in, nodes_out)
        self.slave.forward(is_train, self.snodes_in, self.snodes_out)
        for i, (s, n) in enumerate(zip(self.snodes_out, nodes_out)):
            self._cmp("forward.out[%d]" % i, n.data, s.data)

    def backprop(self, prop_grad, nodes_in, nodes_out):
        if not self.master.self_loop:
            for s, n in zip(self.snodes_out, nodes_out):
                s.data = n.data.clone()
        else:
            for s, n in zip(self.snodes_in, nodes_in):
                s.data = n.data.clone()
        for (_, _, mg), (_, _, sg) in zip(self.master.weights(), self.slave.weights()):
            sg.copy_(mg.reshape(sg.shape))
        self.master.backprop(prop_grad, nodes_in, nodes_out)
        self.slave.backprop(prop_grad, self.snodes_in, self.snodes_out)
        for (tag, _, mg), (_, _, sg) in zip(self.master.weights(), self.slave.weights()):
            self._cmp("backprop.grad[%s]" % tag, mg, sg)
        if prop_grad:
            for i, (s, n) in enumerate(zip(self.snodes_in, nodes_in)):
                self._cmp("backprop.in[%d]" % i, n.data, s.data)
