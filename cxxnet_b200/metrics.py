"""Evaluation metrics: error, rmse, logloss, rec@n with the reference's print format.

Parity with reference ``src/utils/metric.h:20-236``: ``rmse`` is the per-instance *sum of squared
error* (no root, ``:73-89``); ``logloss`` clamps to 1e-15; ``rec@n`` breaks score ties
randomly; ``MetricSet.print`` emits ``\\t<data>-<metric>[<field>]:<value>``.

Unlike the reference (host loops over a D2H copy every step, K20 in SURVEY.md section 2.9) the
statistics here are torch reductions on whatever device the scores live on and are kept
in device scalars; the host reads them only when a line is printed.
"""
from __future__ import annotations

import re
from typing import Dict, List

import torch


class Metric:
    name = "metric"

    def __init__(self):
        self.clear()

    def clear(self) -> None:
        # the accumulator is a persistent device scalar updated in place: a training step captured into
        # a CUDA graph keeps adding to the same address on replay
        if getattr(self, "sum_metric", None) is not None:
            self.sum_metric.zero_()
        else:
            self.sum_metric = None  # device scalar (float64), created by the first add_eval
        self.cnt_inst = 0

    def add_eval(self, pred: torch.Tensor, label: torch.Tensor) -> None:
        """pred [n, k] scores, label [n, w] label field."""
        if pred.shape[0] == 0:
            return
        v = self.calc(pred.float(), label.to(pred.device).float()).double().sum()
        if self.sum_metric is None or self.sum_metric.device != v.device:
            self.sum_metric = v.clone()
        else:
            self.sum_metric.add_(v)
        self.cnt_inst += int(pred.shape[0])

    def calc(self, pred, label) -> torch.Tensor:
        raise NotImplementedError

    def get(self) -> float:
        if self.cnt_inst == 0 or self.sum_metric is None:
            return float("nan")
        return float(self.sum_metric.item()) / self.cnt_inst


class MetricRMSE(Metric):
    name = "rmse"

    def calc(self, pred, label):
        if pred.shape[1] != label.shape[1]:
            raise ValueError("Metric: In RMSE metric, the size of prediction and label must be same.")
        return ((pred - label) ** 2).sum(1)


class MetricError(Metric):
    name = "error"

    def calc(self, pred, label):
        if pred.shape[1] != 1:
            # first maximum wins, as the reference's strict '>' scan
            idx = pred.argmax(dim=1)
            mx = pred.max(dim=1, keepdim=True).values
            first = (pred == mx).float().argmax(dim=1)
            idx = first
        else:
            idx = (pred[:, 0] > 0.0).long()
        return (idx != label[:, 0].long()).float()


class MetricLogloss(Metric):
    name = "logloss"

    def calc(self, pred, label):
        if pred.shape[1] != 1:
            tgt = label[:, 0].long().clamp(0, pred.shape[1] - 1)
            p = pred.gather(1, tgt.view(-1, 1))[:, 0]
            return -torch.log(p.clamp(1e-15, 1.0 - 1e-15))
        py = pred[:, 0].clamp(1e-15, 1.0 - 1e-15)
        y = label[:, 0]
        res = -(y * torch.log(py) + (1.0 - y) * torch.log(1 - py))
        return res


class MetricRecall(Metric):
    def __init__(self, name: str):
        m = re.match(r"rec@(\d+)", name)
        if not m:
            raise ValueError("must specify n for rec@n")
        self.topn = int(m.group(1))
        self.name = name
        super().__init__()

    def calc(self, pred, label):
        if pred.shape[1] < self.topn:
            raise ValueError("it is meaningless to take rec@n for list shorter than n, evaluating rec@%d, list=%d"
                             % (self.topn, pred.shape[1]))
        # random tie-break: add a tiny random permutation order as secondary key
        noise = torch.rand(pred.shape, device=pred.device)
        order = torch.argsort(noise, dim=1)
        shuffled = pred.gather(1, order)
        top = shuffled.topk(self.topn, dim=1).indices
        top_idx = order.gather(1, top)  # [n, topn] original class ids
        lab = label.long()  # [n, w]
        hit = (top_idx.unsqueeze(2) == lab.unsqueeze(1)).any(dim=2).float().sum(1)
        return hit / label.shape[1]


def create_metric(name: str) -> Metric:
    if name == "rmse":
        return MetricRMSE()
    if name == "error":
        return MetricError()
    if name == "logloss":
        return MetricLogloss()
    if name.startswith("rec@"):
        return MetricRecall(name)
    raise ValueError("Metric: Unknown metric name: %s\n" % name)


class MetricSet:
    def __init__(self):
        self.evals: List[Metric] = []
        self.label_fields: List[str] = []

    def add_metric(self, name: str, field: str) -> None:
        self.evals.append(create_metric(name))
        self.label_fields.append(field)

    def clear(self) -> None:
        for e in self.evals:
            e.clear()

    def add_eval(self, predscores: List[torch.Tensor], fields: List[torch.Tensor], name2findex: Dict[str, int]):
        assert len(predscores) == len(self.evals), \
            "Metric: Number of predict scores and number of metrics should be equal."
        for e, f, p in zip(self.evals, self.label_fields, predscores):
            if f not in name2findex:
                raise ValueError("Metric: unknown target = %s" % f)
            e.add_eval(p, fields[name2findex[f]])

    def print(self, evname: str) -> str:
        out = []
        for e, f in zip(self.evals, self.label_fields):
            s = "\t%s-%s" % (evname, e.name)
            if f != "label":
                s += "[%s]" % f
            s += ":%g" % e.get()
            out.append(s)
        return "".join(out)
