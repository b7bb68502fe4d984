"""Weight updaters: sgd / nag / adam with cxxnet's exact formulas and learning-rate schedules.

``UpdaterParam`` mirrors reference ``src/updater/param.h:13-133`` (defaults, tag-scoped keys such as
``wmat:lr`` / ``bias:wd``, ``lr:schedule`` family).  "epoch" is the number of updates done so
far (mini-batches), not passes over the data (reference ``updater.h:48-51``).

The classes below are the *reference backend* updaters (plain torch, one per visited
``(tag, w, dw)``); the b200 backend evaluates the same ``UpdaterParam`` schedules on the host
and applies them with one multi-tensor kernel (``parallel/arena.py``).
"""
from __future__ import annotations

import math
import sys
from typing import List

import numpy as np
import torch


def _f32(x: float) -> float:
    return float(np.float32(x))


class UpdaterParam:
    def __init__(self, tag: str = ""):
        self.tag = tag
        self.round = 0
        self.silent = 0
        self.learning_rate = 0.01
        self.wd = 0.0
        self.momentum = 0.9
        self.lr_schedule = 0
        self.momentum_schedule = 0
        self.base_lr = 0.01
        self.lr_step = 1
        self.lr_gamma = 0.5
        self.lr_alpha = 0.5
        self.lr_factor = 0.1
        self.lr_minimum = 0.00001
        self.start_epoch = 0
        self.base_momentum = 0.5
        self.final_momentum = 0.90
        self.saturation_epoch = 0
        self.clip_gradient = 0.0

    def schedule_epoch(self, epoch: int) -> None:
        """reference ``param.h:76-94``; note expdecay divides as float, polydecay / factor as integers."""
        if self.lr_schedule == 0:
            lr = self.base_lr
        elif self.lr_schedule == 1:
            lr = self.base_lr * math.pow(self.lr_gamma, float(epoch) / self.lr_step)
        elif self.lr_schedule == 2:
            lr = self.base_lr * math.pow(1.0 + (epoch // self.lr_step) * self.lr_gamma, -self.lr_alpha)
        elif self.lr_schedule == 3:
            lr = self.base_lr * math.pow(self.lr_factor, epoch // self.lr_step)
        else:
            raise ValueError("unknown schedule type")
        if self.momentum_schedule and self.saturation_epoch:
            self.momentum += (self.final_momentum - self.base_momentum) / self.saturation_epoch * epoch \
                + self.base_momentum
        self.momentum = self.momentum if self.momentum < self.final_momentum else self.final_momentum
        lr = self.lr_minimum if lr < self.lr_minimum else lr
        if epoch < self.start_epoch:
            lr = self.base_lr
        self.learning_rate = _f32(lr)

    def set_param(self, name: str, val: str) -> None:
        if self.tag and name.startswith(self.tag) and len(name) > len(self.tag) and name[len(self.tag)] == ":":
            name = name[len(self.tag) + 1:]
        if name in ("lr", "eta"): self.base_lr = _f32(float(val))
        if name == "wd": self.wd = _f32(float(val))
        if name == "momentum": self.momentum = _f32(float(val))
        if name == "silent": self.silent = int(val)
        if name == "momentum_schedule": self.momentum_schedule = int(val)
        if name == "clip_gradient": self.clip_gradient = _f32(float(val))
        if name == "final_momentum": self.final_momentum = _f32(float(val))
        if name == "base_momentum": self.base_momentum = _f32(float(val))
        if name == "saturation_epoch": self.saturation_epoch = int(val)
        sub = None
        if name.startswith("lr:"):
            sub = name[3:]
        elif name.startswith("eta:"):
            sub = name[4:]
        if sub is not None:
            if sub == "schedule":
                m = {"constant": 0, "expdecay": 1, "polydecay": 2, "factor": 3}
                if val in m:
                    self.lr_schedule = m[val]
            if sub == "gamma": self.lr_gamma = _f32(float(val))
            if sub == "alpha": self.lr_alpha = _f32(float(val))
            if sub == "step": self.lr_step = int(val)
            if sub == "factor": self.lr_factor = _f32(float(val))
            if sub == "minimum_lr": self.lr_minimum = _f32(float(val))
            if sub == "start_epoch": self.start_epoch = int(val)


class Updater:
    """reference backend updater base (reference ``updater.h:21-68`` IUpdater)."""
    kind = "sgd"

    def __init__(self, w: torch.Tensor, dw: torch.Tensor, tag: str):
        self.w, self.dw = w, dw
        self.param = UpdaterParam(tag)

    def set_param(self, name: str, val: str) -> None:
        self.param.set_param(name, val)

    def init(self) -> None:
        pass

    def start_round(self, r: int) -> None:
        self.param.round = r

    def update(self, epoch: int) -> None:
        self.apply_update(epoch, self.dw)
        self.dw.zero_()

    def apply_update(self, epoch: int, grad: torch.Tensor) -> None:
        raise NotImplementedError


def _clip(g: torch.Tensor, c: float) -> torch.Tensor:
    """reference ``sgd_updater-inl.hpp:15-22``: NaN -> 0, clamp to [-c, c]."""
    return torch.nan_to_num(g, nan=0.0).clamp(-c, c)


class SGDUpdater(Updater):
    """``m = mu*m - lr*(clip?(g) + wd*w); w += m`` (reference ``src/updater/sgd_updater-inl.hpp:26-85``)."""
    kind = "sgd"

    def init(self):
        if self.param.silent == 0:
            print("SGDUpdater: eta=%f, mom=%f" % (self.param.base_lr, self.param.momentum))
        self.m_w = torch.zeros_like(self.w)

    def apply_update(self, epoch, grad):
        p = self.param
        p.schedule_epoch(epoch)
        self.m_w.mul_(p.momentum)
        g = _clip(grad, p.clip_gradient) if p.clip_gradient != 0.0 else grad
        self.m_w.add_(g + p.wd * self.w, alpha=-p.learning_rate)
        self.w.add_(self.m_w)


class NAGUpdater(Updater):
    """``old=m; m = mu*m - lr*(g+wd*w); w += (1+mu)*m - mu*old`` (reference ``src/updater/nag_updater-inl.hpp:65-72``)."""
    kind = "nag"

    def init(self):
        if self.param.silent == 0:
            print("NAGUpdater: eta=%f, mom=%f" % (self.param.base_lr, self.param.momentum))
        self.m_w = torch.zeros_like(self.w)

    def apply_update(self, epoch, grad):
        p = self.param
        p.schedule_epoch(epoch)
        old = self.m_w.clone()
        self.m_w.mul_(p.momentum).add_(grad + p.wd * self.w, alpha=-p.learning_rate)
        self.w.add_((1.0 + p.momentum) * self.m_w - p.momentum * old)


class AdamUpdater(Updater):
    """reference ``src/updater/adam_updater-inl.hpp:19-24,73-81``: ``beta1``/``beta2`` keys are the decay rates
    (1 - beta of the paper, defaults 0.1 / 0.001), weight decay enters with the sign the reference
    uses (``g -= wd*w``), and the lr schedule is ignored (``base_lr`` only)."""
    kind = "adam"

    def __init__(self, w, dw, tag):
        super().__init__(w, dw, tag)
        self.decay1, self.decay2 = 0.1, 0.001

    def set_param(self, name, val):
        super().set_param(name, val)
        if name == "beta1": self.decay1 = float(val)
        if name == "beta2": self.decay2 = float(val)

    def init(self):
        if self.param.silent == 0:
            print("AdamUpdater: eta=%f, beta1=%f, beta2=%f" % (self.param.base_lr, self.decay1, self.decay2))
        self.m_w1 = torch.zeros_like(self.w)
        self.m_w2 = torch.zeros_like(self.w)

    def apply_update(self, epoch, grad):
        p = self.param
        if p.wd > 0.0:
            grad.sub_(p.wd * self.w)
        fix1 = 1.0 - math.pow(1.0 - self.decay1, epoch + 1)
        fix2 = 1.0 - math.pow(1.0 - self.decay2, epoch + 1)
        lr_t = p.base_lr * math.sqrt(fix2) / fix1
        self.m_w1.add_(self.decay1 * (grad - self.m_w1))
        self.m_w2.add_(self.decay2 * (grad * grad - self.m_w2))
        self.w.sub_(lr_t * (self.m_w1 / (torch.sqrt(self.m_w2) + 1e-8)))


def create_updater(kind: str, w: torch.Tensor, dw: torch.Tensor, tag: str) -> Updater:
    """reference ``src/updater/updater_impl-inl.hpp:17-32``."""
    if kind == "sgd":
        return SGDUpdater(w, dw, tag)
    if kind == "nag":
        return NAGUpdater(w, dw, tag)
    if kind == "adam":
        return AdamUpdater(w, dw, tag)
    raise ValueError("unknown updater type %s" % kind)


def encode_data_key(layer_index: int, tag: str) -> int:
    """parameter-server key of a weight (reference ``updater.h:150-173``)."""
    if tag == "bias":
        return layer_index * 4 + 1
    if tag == "wmat":
        return layer_index * 4 + 0
    raise ValueError("EncodeDataKey: only support weight tag: wmat or bias")


def decode_tag(key: int) -> str:
    return {0: "wmat", 1: "bias"}[key % 4]
