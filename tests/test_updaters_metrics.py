"""updaters / schedules (reference src/updater/*.hpp, param.h) and metrics (src/utils/metric.h)."""
import math

import numpy as np
import pytest
import torch

from cxxnet_b200.metrics import MetricSet, create_metric
from cxxnet_b200.updaters import UpdaterParam, create_updater, decode_tag, encode_data_key


def test_schedules():
    p = UpdaterParam("wmat")
    p.set_param("lr", "0.1")
    p.schedule_epoch(7)
    assert p.learning_rate == pytest.approx(0.1)
    p.set_param("lr:schedule", "expdecay"); p.set_param("lr:gamma", "0.1"); p.set_param("lr:step", "100")
    p.schedule_epoch(50)
    assert p.learning_rate == pytest.approx(0.1 * 0.1 ** 0.5, rel=1e-5)      # float division
    p.set_param("lr:schedule", "factor"); p.set_param("lr:factor", "0.5"); p.set_param("lr:step", "10")
    p.schedule_epoch(25)
    assert p.learning_rate == pytest.approx(0.1 * 0.25, rel=1e-6)            # integer division
    p.set_param("lr:schedule", "polydecay"); p.set_param("lr:gamma", "2"); p.set_param("lr:alpha", "0.5")
    p.schedule_epoch(30)
    assert p.learning_rate == pytest.approx(0.1 * (1 + 3 * 2) ** -0.5, rel=1e-6)
    p.set_param("lr:schedule", "factor"); p.set_param("lr:factor", "1e-9")
    p.schedule_epoch(100)
    assert p.learning_rate == pytest.approx(1e-5)                            # lr_minimum clamp
    p.set_param("lr:start_epoch", "1000")
    p.schedule_epoch(100)
    assert p.learning_rate == pytest.approx(0.1)


def test_tag_scoped_keys():
    w, b = UpdaterParam("wmat"), UpdaterParam("bias")
    for p in (w, b):
        for k, v in [("lr", "0.5"), ("wmat:lr", "0.01"), ("bias:lr", "0.02"), ("wmat:wd", "0.0005"), ("bias:wd", "0")]:
            p.set_param(k, v)
    assert w.base_lr == pytest.approx(0.01) and b.base_lr == pytest.approx(0.02)
    assert w.wd == pytest.approx(0.0005) and b.wd == 0.0


def _mk(kind, **kw):
    torch.manual_seed(0)
    w = torch.randn(5, 3)
    g = torch.randn(5, 3)
    u = create_updater(kind, w, g, "wmat")
    u.param.silent = 1
    for k, v in kw.items():
        u.set_param(k, str(v))
    u.init()
    return u, w, g


def test_sgd_formula_and_grad_zeroing():
    u, w, g = _mk("sgd", lr=0.1, momentum=0.9, wd=0.01)
    w0, g0 = w.clone(), g.clone()
    u.update(0)
    m1 = -0.1 * (g0 + 0.01 * w0)
    assert torch.allclose(w, w0 + m1, atol=1e-6) and float(g.abs().max()) == 0.0
    g.copy_(g0)
    w1 = w.clone()
    u.update(1)
    m2 = 0.9 * m1 - 0.1 * (g0 + 0.01 * w1)
    assert torch.allclose(w, w1 + m2, atol=1e-6)


def test_sgd_clip_maps_nan_to_zero():
    u, w, g = _mk("sgd", lr=1.0, momentum=0.0, clip_gradient=0.5)
    g.fill_(3.0); g[0, 0] = float("nan")
    w0 = w.clone()
    u.update(0)
    assert torch.isfinite(w).all()
    assert w[0, 0] == pytest.approx(float(w0[0, 0])) and w[1, 1] == pytest.approx(float(w0[1, 1]) - 0.5)


def test_nag_formula():
    u, w, g = _mk("nag", lr=0.1, momentum=0.9)
    w0, g0 = w.clone(), g.clone()
    u.update(0)
    m1 = -0.1 * g0
    assert torch.allclose(w, w0 + 1.9 * m1, atol=1e-6)


def test_adam_reference_quirks():
    u, w, g = _mk("adam", lr=0.001, wd=0.1)
    w0, g0 = w.clone(), g.clone()
    u.update(0)
    gr = g0 - 0.1 * w0                      # sign as written in the reference
    m1, m2 = 0.1 * gr, 0.001 * gr * gr
    lr_t = 0.001 * math.sqrt(1 - 0.999) / (1 - 0.9)
    assert torch.allclose(w, w0 - lr_t * m1 / (m2.sqrt() + 1e-8), atol=1e-6)


def test_ps_key_encoding():
    assert encode_data_key(3, "wmat") == 12 and encode_data_key(3, "bias") == 13
    assert decode_tag(12) == "wmat" and decode_tag(13) == "bias"


def test_metrics_values_and_print_format():
    pred = torch.tensor([[0.1, 0.7, 0.2], [0.5, 0.25, 0.25], [0.2, 0.2, 0.6]])
    label = torch.tensor([[1.0], [2.0], [2.0]])
    ms = MetricSet()
    for name in ("error", "logloss", "rec@2"):
        ms.add_metric(name, "label")
    ms.add_eval([pred, pred, pred], [label], {"label": 0})
    err, ll, rec = [e.get() for e in ms.evals]
    assert err == pytest.approx(1 / 3)
    assert ll == pytest.approx(-(math.log(0.7) + math.log(0.25) + math.log(0.6)) / 3, rel=1e-5)
    assert rec == pytest.approx(1.0)      # label 2 ties with 1 in row 1: top-2 of {0.5,0.25,0.25} always holds one of them
    s = ms.print("test")
    assert s.startswith("\ttest-error:") and "\ttest-logloss:" in s and "\ttest-rec@2:" in s


def test_rmse_is_sum_of_squares_and_binary_error():
    m = create_metric("rmse")
    m.add_eval(torch.tensor([[1.0, 2.0], [0.0, 0.0]]), torch.tensor([[0.0, 0.0], [1.0, 1.0]]))
    assert m.get() == pytest.approx((5.0 + 2.0) / 2)          # no root (reference metric.h:73-89)
    e = create_metric("error")
    e.add_eval(torch.tensor([[0.3], [-0.2]]), torch.tensor([[1.0], [1.0]]))
    assert e.get() == pytest.approx(0.5)
    with pytest.raises(ValueError):
        create_metric("bogus")


def test_metric_field_name_in_print():
    ms = MetricSet()
    ms.add_metric("error", "aux")
    ms.add_eval([torch.tensor([[0.9, 0.1]])], [torch.zeros(1, 1), torch.zeros(1, 1)], {"label": 0, "aux": 1})
    assert ms.print("train") == "\ttrain-error[aux]:0"
