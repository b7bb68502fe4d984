"""Shared helpers: build trainers from conf text, run the b200 backend against the reference
backend ("pairtest" at net level)."""
import numpy as np
import torch

from cxxnet_b200.config import parse_config_string
from cxxnet_b200.io.data import DataBatch
from cxxnet_b200.trainer import NetTrainer


def make_trainer(conf: str, **over) -> NetTrainer:
    tr = NetTrainer()
    for k, v in parse_config_string(conf):
        tr.set_param(k, v)
    for k, v in over.items():
        tr.set_param(k, str(v))
    tr.set_param("silent", "1")
    tr.init_model()
    return tr


def make_batch(shape, nclass=10, label_width=1, seed=0, dense_label=None, pin=False) -> DataBatch:
    g = torch.Generator().manual_seed(seed)
    b = DataBatch()
    b.data = torch.randn(shape, generator=g)
    if dense_label is not None:
        b.label = torch.rand((shape[0], dense_label), generator=g)
    else:
        b.label = torch.randint(0, nclass, (shape[0], label_width), generator=g).float()
    if pin and torch.cuda.is_available():
        b.data, b.label = b.data.pin_memory(), b.label.pin_memory()
    b.batch_size = shape[0]
    b.inst_index = np.arange(shape[0], dtype=np.uint32)
    return b


def rel_err(a: torch.Tensor, b: torch.Tensor, floor: float = 0.0) -> float:
    """relative L1 error sum|a-b| / max(sum|a|, floor*n) (the reference pairtest metric, with an
    absolute floor so that quantities that are analytically zero do not blow up)."""
    a, b = a.detach().float().cpu().reshape(-1), b.detach().float().cpu().reshape(-1)
    denom = max(float(a.abs().sum()), floor * a.numel())
    return float((a - b).abs().sum()) / denom if denom > 0 else float((a - b).abs().sum())


def copy_weights(src: NetTrainer, dst: NetTrainer) -> None:
    for i, c in enumerate(src.net.connections):
        for tag, w, _ in c.layer.weights():
            ws, _ = src.net.get_weight(i, tag)
            dst.net.set_weight(i, ws, tag)


def compare_backends(conf: str, batch: DataBatch, tol_out=2e-2, tol_grad=4e-2, check_nodes=None, skip=(),
                     return_trainers=False):
    """one forward/backward on the reference backend (fp32, CUDA) and on the b200 backend with
    identical weights and data; returns dict of relative L1 errors and asserts tolerances."""
    ref = make_trainer(conf, dev="gpu:0", backend="ref")
    b2 = make_trainer(conf, dev="gpu:0", backend="b200")
    copy_weights(ref, b2)
    errs = {}
    for tr in (ref, b2):
        net = tr.net
        lo, hi = tr._slice(batch.batch_size)
        net.label_info.fields = tr._label_fields(batch.label, lo, hi, tr.device)
        net.accumulate_grads = False
        net.set_input(batch.data[lo:hi], [])
        net.forward(True)
    nodes = check_nodes if check_nodes is not None else [len(ref.net.nodes) - 1]
    for nid in nodes:
        errs["node%d" % nid] = rel_err(ref.net.copy_node_data(nid), b2.net.copy_node_data(nid))
    for tr in (ref, b2):
        tr.net.backprop(True, False, 0)
    torch.cuda.synchronize()
    for i, c in enumerate(ref.net.connections):
        for tag, w, g in c.layer.weights():
            errs["grad[%d:%s]" % (i, tag)] = rel_err(ref.net.get_grad(i, tag), b2.net.get_grad(i, tag), floor=1e-4)
    # input gradient
    gin_ref = ref.net.nodes[0].data
    gin_b2 = b2.net._node_load(b2.net.nodes[0], grad=True)
    errs["grad_in"] = rel_err(gin_ref, gin_b2)
    for k, v in errs.items():
        if k in skip:
            continue
        tol = tol_out if k.startswith("node") else tol_grad
        assert v == v and v <= tol, "%s: relative error %g > %g (all: %s)" % (k, v, tol, errs)
    if return_trainers:
        return errs, ref, b2
    return errs
