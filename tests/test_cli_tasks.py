"""task driver: train / continue / pred / extract / finetune, model naming, metric line format
(reference src/cxxnet_main.cpp), plus pairtest and weight get/set through the trainer."""
import os
import re
import subprocess
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CONF = """
data = train
iter = mnist
    path_img = "none"
    path_label = "none"
    num_synthetic = 3000
    shuffle = 1
iter = end
eval = test
iter = mnist
    path_img = "none"
    path_label = "none"
    num_synthetic = 200
    index_offset = 7
iter = end
pred = %(pred)s
iter = mnist
    path_img = "none"
    path_label = "none"
    num_synthetic = 200
    index_offset = 7
iter = end
netconfig=start
layer[+1:fc1] = fullc:fc1
  nhidden = 32
  init_sigma = 0.01
layer[+1:sg1] = sigmoid:se1
layer[sg1->fc2] = fullc:fc2
  nhidden = 10
  init_sigma = 0.01
layer[+0] = softmax
netconfig=end
input_shape = 1,1,784
batch_size = 100
dev = cpu
save_model = 1
num_round = %(rounds)d
max_round = %(rounds)d
eta = 0.1
momentum = 0.9
metric[label] = error
model_dir = %(model_dir)s
silent = 1
"""


def _run(args, cwd):
    env = dict(os.environ)
    env["PYTHONPATH"] = ROOT + os.pathsep + env.get("PYTHONPATH", "")
    r = subprocess.run([sys.executable, "-m", "cxxnet_b200"] + args, cwd=cwd, env=env, stdout=subprocess.PIPE,
                       stderr=subprocess.PIPE, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    return r


def test_train_continue_pred_extract_finetune(tmp_path):
    mdir = tmp_path / "models"
    conf = tmp_path / "mnist.conf"
    conf.write_text(CONF % {"rounds": 2, "model_dir": mdir, "pred": tmp_path / "pred.txt"})
    r = _run([str(conf)], tmp_path)
    assert sorted(os.listdir(mdir)) == ["0000.model", "0001.model", "0002.model"]
    lines = [l for l in r.stderr.splitlines() if l.startswith("[")]
    assert len(lines) == 2
    m = re.match(r"\[1\]\ttrain-error:([0-9.eE+-]+)\ttest-error:([0-9.eE+-]+)$", lines[0])
    assert m, lines[0]
    assert float(re.match(r"\[2\]\ttrain-error:([0-9.eE+-]+)", lines[1]).group(1)) < min(float(m.group(1)), 0.2)
    # continue=1 picks up after the latest model and trains the remaining rounds
    r2 = _run([str(conf), "continue=1", "num_round=3", "max_round=3"], tmp_path)
    assert "Continue training from round 3" in r2.stdout
    assert "0003.model" in os.listdir(mdir)
    assert [l for l in r2.stderr.splitlines() if l.startswith("[3]")]
    # pred: one class id per line, as many as instances
    _run([str(conf), "task=pred", "model_in=%s" % (mdir / "0003.model")], tmp_path)
    preds = (tmp_path / "pred.txt").read_text().split()
    assert len(preds) == 200 and all(p.isdigit() and 0 <= int(p) <= 9 for p in preds)
    # extract: text rows + meta "nrow,c,h,w"; top[-1] is the softmax output
    _run([str(conf), "task=extract", "model_in=%s" % (mdir / "0003.model"), "extract_node_name=top[-1]",
          "output_format=txt"], tmp_path)
    rows = (tmp_path / "pred.txt").read_text().strip().split("\n")
    assert len(rows) == 200 and len(rows[0].split()) == 10
    assert abs(sum(float(v) for v in rows[0].split()) - 1.0) < 1e-4
    assert (tmp_path / "pred.txt.meta").read_text().strip() == "200,1,1,10"
    _run([str(conf), "task=extract", "model_in=%s" % (mdir / "0003.model"), "extract_node_name=sg1",
          "output_format=bin"], tmp_path)
    assert os.path.getsize(tmp_path / "pred.txt") == 200 * 32 * 4
    # pred_raw (documented by the reference, dispatched here): one text row of raw outputs per instance
    _run([str(conf), "task=pred_raw", "model_in=%s" % (mdir / "0003.model")], tmp_path)
    raw = (tmp_path / "pred.txt").read_text().strip().split("\n")
    assert len(raw) == 200 and len(raw[0].split()) == 10
    assert [int(np.argmax([float(v) for v in ln.split()])) for ln in raw] == [int(p) for p in preds]
    # finetune: layers with matching names are copied, epoch counter restarts
    ft = tmp_path / "ft.conf"
    ft.write_text((CONF % {"rounds": 1, "model_dir": tmp_path / "ft_models", "pred": tmp_path / "p2.txt"})
                  .replace("fullc:fc2", "fullc:newhead"))
    r3 = _run([str(ft), "task=finetune", "model_in=%s" % (mdir / "0003.model")], tmp_path)
    assert "Copying layer fc1" in r3.stdout and "Copying layer newhead" not in r3.stdout


def test_test_io_mode_runs_pipeline_only(tmp_path):
    conf = tmp_path / "c.conf"
    conf.write_text(CONF % {"rounds": 1, "model_dir": tmp_path / "m", "pred": tmp_path / "p.txt"})
    r = _run([str(conf), "test_io=1", "silent=0"], tmp_path)
    assert "start I/O test" in r.stdout and "train-error" not in r.stderr
    assert re.search(r"I/O test round 0: 30 batches, 3000 instances in [0-9.]+ sec: \d+ instances/sec", r.stdout)


def test_pairtest_layer_reports_zero_error_for_identical_implementations(capsys):
    from util import make_batch, make_trainer
    conf = """
netconfig=start
layer[+1] = pairtest-conv-conv:pc
  kernel_size = 3
  nchannel = 4
  pad = 1
layer[+1] = pairtest-relu-relu
layer[+1] = flatten
layer[+1] = pairtest-fullc-fullc:pf
  nhidden = 5
layer[+0] = softmax
netconfig=end
input_shape = 2,6,6
batch_size = 4
dev = cpu
eta = 0.1
metric = error
"""
    tr = make_trainer(conf)
    for s in range(2):
        tr.update(make_batch((4, 2, 6, 6), nclass=5, seed=s))
    reports = []
    for c in tr.net.connections:
        reports += getattr(c.layer, "reports", [])
    assert len(reports) >= 10 and max(r[1] for r in reports) <= 1e-6
    assert "PairTest" not in capsys.readouterr().err
    # a deliberately different slave is reported (never fails the run)
    tr2 = make_trainer(conf.replace("pairtest-relu-relu", "pairtest-relu-tanh"))
    tr2.update(make_batch((4, 2, 6, 6), nclass=5))
    assert "PairTest[relu-tanh]" in capsys.readouterr().err


def test_get_set_weight_roundtrip_and_update_period():
    from cxxnet_b200.models import mlp_conf
    from util import make_batch, make_trainer
    tr = make_trainer(mlp_conf(dev="cpu"))
    w, shape = tr.get_weight("fc1", "wmat")
    assert shape == [100, 784] and w.shape == (100, 784)
    tr.set_weight(np.zeros_like(w), "fc1", "wmat")
    assert float(np.abs(tr.get_weight("fc1", "wmat")[0]).max()) == 0.0
    tr.set_weight(w, "fc1", "wmat")
    assert np.array_equal(tr.get_weight("fc1", "wmat")[0], w)
    with pytest.raises(ValueError):
        tr.get_weight("fc1", "gamma")
    # update_period=2: weights move only every second call and equal one step on the summed gradient
    a = make_trainer(mlp_conf(dev="cpu", batch_size=50, extra="update_period = 2\nmomentum = 0\n"))
    b = make_trainer(mlp_conf(dev="cpu", batch_size=100, extra="momentum = 0\n"))
    for name in ("fc1", "fc2"):
        for tag in ("wmat", "bias"):
            b.set_weight(a.get_weight(name, tag)[0], name, tag)
    full = make_batch((100, 1, 1, 784), seed=5)
    from cxxnet_b200.io.data import DataBatch
    halves = []
    for lo in (0, 50):
        h = DataBatch()
        h.data, h.label, h.batch_size = full.data[lo:lo + 50], full.label[lo:lo + 50], 50
        halves.append(h)
    w0 = a.get_weight("fc2", "wmat")[0].copy()
    a.update(halves[0])
    assert np.array_equal(a.get_weight("fc2", "wmat")[0], w0) and a.epoch_counter == 0
    a.update(halves[1])
    b.update(full)
    assert a.epoch_counter == 1
    assert np.allclose(a.get_weight("fc2", "wmat")[0], b.get_weight("fc2", "wmat")[0], atol=1e-6)


def test_layer_timing_reports_every_layer(tmp_path):
    """`layer_timing = 1`: a forward / backward table per round on stdout, one row per layer"""
    conf = tmp_path / "mnist.conf"
    conf.write_text(CONF % {"rounds": 1, "model_dir": tmp_path / "models", "pred": tmp_path / "pred.txt"})
    r = _run([str(conf), "layer_timing=1"], tmp_path)
    assert "layer timing (mean us per step" in r.stdout
    rows = [l for l in r.stdout.splitlines() if re.match(r"\s+\d+ (fullc|sigmoid|softmax):", l)]
    assert len(rows) == 4, r.stdout
    fwd = [float(l.split()[2]) for l in rows]
    bwd = [float(l.split()[3]) for l in rows]
    assert all(v > 0 for v in fwd) and all(v > 0 for v in bwd)


CONV_CONF = """
data = train
iter = mnist
    path_img = "none"
    path_label = "none"
    num_synthetic = 300
    input_flat = 0
iter = end
netconfig=start
layer[+1] = conv:cv1
  kernel_size = 3
  stride = 2
  nchannel = 4
layer[+1] = max_pooling
  kernel_size = 3
  stride = 2
layer[+1] = flatten
layer[+1] = fullc:%(head)s
  nhidden = 10
layer[+0] = softmax
netconfig=end
input_shape = 1,28,28
batch_size = 100
dev = cpu
num_round = 1
max_round = 1
eta = 0.1
metric = error
model_dir = %(model_dir)s
silent = 1
"""


def test_finetune_and_model_info_on_a_net_with_weightless_layers(tmp_path):
    """a stored net does not carry its pooling windows: copying layers by name and inspecting the file must not
    need them"""
    a = tmp_path / "a.conf"
    a.write_text(CONV_CONF % {"head": "fc", "model_dir": tmp_path / "m"})
    _run([str(a)], tmp_path)
    model = tmp_path / "m" / "0001.model"
    b = tmp_path / "b.conf"
    b.write_text(CONV_CONF % {"head": "newhead", "model_dir": tmp_path / "m2"})
    r = _run([str(b), "task=finetune", "model_in=%s" % model], tmp_path)
    assert "Copying layer cv1" in r.stdout and "Copying layer newhead" not in r.stdout
    env = dict(os.environ)
    env["PYTHONPATH"] = ROOT + os.pathsep + env.get("PYTHONPATH", "")
    info = subprocess.run([sys.executable, "-m", "cxxnet_b200.tools.model_info", str(model), "--weights"], env=env,
                          stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=120)
    assert info.returncode == 0, info.stdout
    assert "input 1x28x28" in info.stdout and "MaxPooling" in info.stdout and "wmat  (1, 4, 9)" in info.stdout
    assert "parameters: %d" % (4 * 9 + 4 + 10 * 4 * 6 * 6 + 10) in info.stdout


def test_check_nan_names_the_layer():
    from cxxnet_b200.trainer import NetTrainer
    from cxxnet_b200.config import parse_config_string
    from cxxnet_b200.models import mlp_conf
    tr = NetTrainer()
    for k, v in parse_config_string(mlp_conf(dev="cpu", batch_size=10)):
        tr.set_param(k, v)
    tr.set_param("check_nan", "1")
    tr.set_param("silent", "1")
    tr.init_model()
    tr.start_round(1)                                         # fine
    assert tr.find_non_finite() is None
    w, _ = tr.get_weight("fc2", "wmat")
    w = w.copy()
    w.reshape(-1)[3] = float("nan")
    tr.set_weight(w, "fc2", "wmat")
    assert tr.find_non_finite() == ("fc2", "wmat")
    with pytest.raises(RuntimeError, match="fc2"):
        tr.start_round(2)
